"""Emits lowered (host-dialect) Moose computations for the replicated protocols on the predictor path, and the
networking pass that turns cross-host edges into Send / Receive pairs.

What the reference's compiler does for these protocols (the lowering pass expands every ReplicatedPlacement kernel
into the HostPlacement operations of its body, moose/src/compilation/lowering.rs:4-6; the networking pass inserts a
Send/Receive pair per (value, consuming host), moose/src/compilation/networking.rs:75-118) restated as a graph
builder: each method below mirrors one replicated / additive kernel of the reference and places every host
operation on the player the reference places it on.  The resulting graphs have the same shape as the reference's
compiled graphs (tutorials/dotprod-compiled.moose: same operator histogram for the same computation, see
tests/test_lowering.py) and run unchanged on the replay executor -- all parties on one GPU, or one party per GPU with
NCCL send/recv (moose_b200/party.py).  Tensors too large for a literal enter through Input operations.

Share layout: a replicated tensor is [[x00, x10], [x11, x21], [x22, x02]] (replicated/mod.rs:75-77), names of the
operations that produced the shares; shares[i][*] live on player i.  An additive tensor is [a0, a1] on players 0, 1
with player 2 as the provider (the (rep0, rep1) branch of additive/convert.rs:28-37).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np

from .textual import Operation

RING_TY = {64: "HostRing64Tensor", 128: "HostRing128Tensor"}
BIT_TY = "HostBitTensor"


class Lowering:
    def __init__(self, bits: int = 64, players: Sequence[str] = ("player0", "player1", "player2"), sync_seed: int = 0):
        self.bits, self.players = bits, list(players)
        self.ring_ty = RING_TY[bits]
        self.ops: List[Operation] = []
        self.ty: Dict[str, str] = {}
        self.host: Dict[str, int] = {}
        self._rng = np.random.default_rng(sync_seed)  # SyncKey::random() at compile time (zero_share.rs:82-85)
        self._keys = None

    # ---- one host operation -------------------------------------------------------------------------------
    def emit(self, kind: str, h: int, inputs: Sequence[str], out_ty: str, attrs: Optional[Dict[str, str]] = None,
             variadic: bool = False) -> str:
        name = f"op_{len(self.ops)}"
        in_types = [self.ty[i] for i in inputs]
        if variadic:
            in_types = in_types[:1]
        self.ops.append(Operation(name, kind, dict(attrs or {}), in_types, variadic, out_ty, list(inputs), "Host",
                                  [self.players[h]]))
        self.ty[name], self.host[name] = out_ty, h
        return name

    def sync_key(self) -> str:
        return self._rng.bytes(16).hex()

    def _bin(self, kind, h, a, b):
        return self.emit(kind, h, [a, b], self.ty[a])

    def add(self, h, a, b): return self._bin("Add", h, a, b)
    def sub(self, h, a, b): return self._bin("Sub", h, a, b)
    def mul(self, h, a, b): return self._bin("Mul", h, a, b)
    def dot(self, h, a, b): return self._bin("Dot", h, a, b)
    def xor(self, h, a, b): return self._bin("Xor", h, a, b)
    def and_(self, h, a, b): return self._bin("And", h, a, b)
    def neg(self, h, a): return self.emit("Neg", h, [a], self.ty[a])
    def shl(self, h, amount, a): return self.emit("Shl", h, [a], self.ty[a], {"amount": str(amount)})
    def shr(self, h, amount, a): return self.emit("Shr", h, [a], self.ty[a], {"amount": str(amount)})
    def shape(self, h, a): return self.emit("Shape", h, [a], "HostShape")
    def key_gen(self, h): return self.emit("PrfKeyGen", h, [], "HostPrfKey")
    def derive_seed(self, h, key, sync): return self.emit("DeriveSeed", h, [key], "HostSeed", {"sync_key": sync})

    def fill(self, h, shape, value: int, ty: str):
        lit = {RING_TY[64]: "Ring64", RING_TY[128]: "Ring128", BIT_TY: "Bit"}[ty]
        return self.emit("Fill", h, [shape], ty, {"value": f"{lit}({value})"})

    def sample_seeded(self, h, shape, seed, ty, max_value=None):
        return self.emit("SampleSeeded", h, [shape, seed], ty, {} if max_value is None else {"max_value": str(max_value)})

    def sample(self, h, shape, ty, max_value=None):
        return self.emit("Sample", h, [shape], ty, {} if max_value is None else {"max_value": str(max_value)})

    def bit_decompose(self, h, a): return self.emit("BitDecompose", h, [a], BIT_TY)
    def ring_inject(self, h, a, bit_idx=0): return self.emit("RingInject", h, [a], self.ring_ty, {"bit_idx": str(bit_idx)})

    def shl_dim(self, h, a, amount, bit_length):
        return self.emit("ShlDim", h, [a], BIT_TY, {"amount": str(amount), "bit_length": str(bit_length)})

    def index_axis(self, h, a, axis, index):
        return self.emit("IndexAxis", h, [a], self.ty[a], {"axis": str(axis), "index": str(index)})

    def input(self, h, arg_name: str, ty: str = "HostFloat64Tensor"):
        return self.emit("Input", h, [], ty, {"arg_name": f'"{arg_name}"'})

    def constant(self, h, values) -> str:
        return self.emit("Constant", h, [], "HostFloat64Tensor", {"value": f"HostFloat64Tensor({np.asarray(values, dtype=np.float64).tolist()})"})

    def fixed_encode(self, h, a, frac: int):
        return self.emit("RingFixedpointEncode", h, [a], self.ring_ty, {"scaling_base": "2", "scaling_exp": str(frac)})

    def fixed_decode(self, h, a, frac: int):
        return self.emit("RingFixedpointDecode", h, [a], "HostFloat64Tensor", {"scaling_base": "2", "scaling_exp": str(frac)})

    def output(self, h, a, tag: str):
        return self.emit("Output", h, [a], self.ty[a], {"tag": f'"{tag}"'})

    # ---- setup, seeds, zero shares (replicated/setup.rs:39-58, zero_share.rs:82-156) ------------------------
    def setup(self):
        if self._keys is None:
            k0, k1, k2 = self.key_gen(0), self.key_gen(1), self.key_gen(2)
            self._keys = [[k0, k1], [k1, k2], [k2, k0]]
        return self._keys

    def gen_seeds(self):
        (k00, k10), (k11, k21), (k22, k02) = self.setup()
        sk0, sk1, sk2 = self.sync_key(), self.sync_key(), self.sync_key()
        s00, s10 = self.derive_seed(0, k00, sk0), self.derive_seed(0, k10, sk1)
        s11, s21 = self.derive_seed(1, k11, sk1), self.derive_seed(1, k21, sk2)
        s22, s02 = self.derive_seed(2, k22, sk2), self.derive_seed(2, k02, sk0)
        return [[s00, s10], [s11, s21], [s22, s02]]

    def zero_share(self, shapes, ty):
        seeds = self.gen_seeds()
        alphas = []
        for i in range(3):
            r0 = self.sample_seeded(i, shapes[i], seeds[i][0], ty)
            r1 = self.sample_seeded(i, shapes[i], seeds[i][1], ty)
            alphas.append(self.sub(i, r0, r1))
        return alphas

    # ---- share / reveal (replicated/convert.rs:49-160, 280-313) --------------------------------------------
    def rep_share(self, x: str, owner: int):
        keys = self.setup()
        o, nxt, prv = owner, (owner + 1) % 3, (owner + 2) % 3
        ty = self.ty[x]
        sk = self.sync_key()
        shape = self.shape(o, x)
        seed_o = self.derive_seed(o, keys[o][0], sk)
        x_oo = self.sample_seeded(o, shape, seed_o, ty)           # x_o at the owner
        x_next = self.sub(o, x, x_oo)                             # x_{o+1} at the owner (and, moved, at o+1)
        seed_p = self.derive_seed(prv, keys[prv][1], sk)          # the same key, held by player o-1
        zero_p = self.fill(prv, shape, 0, ty)
        x_op = self.sample_seeded(prv, shape, seed_p, ty)         # x_o again at player o-1
        zero_n = self.fill(nxt, shape, 0, ty)
        shares = [None, None, None]
        shares[o] = [x_oo, x_next]
        shares[nxt] = [x_next, zero_n]
        shares[prv] = [zero_p, x_op]
        return shares

    def rep_reveal(self, x, receiver: int):
        (x00, x10), (x11, x21), (x22, x02) = x
        if receiver == 0:
            return self.add(0, self.add(0, x00, x10), x21)
        if receiver == 1:
            return self.add(1, self.add(1, x02, x11), x21)
        return self.add(2, self.add(2, x02, x10), x22)

    # ---- share-wise ops (replicated/arith.rs:7-121, 197-315, 525-609) ----------------------------------------
    def _sharewise(self, f, *xs):
        return [[f(i, *[x[i][j] for x in xs]) for j in range(2)] for i in range(3)]

    def rep_add(self, x, y): return self._sharewise(self.add, x, y)
    def rep_sub(self, x, y): return self._sharewise(self.sub, x, y)
    def rep_xor(self, x, y): return self._sharewise(self.xor, x, y)
    def rep_neg(self, x): return self._sharewise(self.neg, x)
    def rep_shl(self, amount, x): return self._sharewise(lambda i, a: self.shl(i, amount, a), x)
    def rep_shl_dim(self, amount, bit_length, x): return self._sharewise(lambda i, a: self.shl_dim(i, a, amount, bit_length), x)
    def rep_index(self, index, x): return self._sharewise(lambda i, a: self.index_axis(i, a, 0, index), x)

    def rep_add_mir(self, x, m):
        """AddOp::rep_mir_kernel (arith.rs:81-121): the public value lands on share 0 (x00 and x02)."""
        (x00, x10), (x11, x21), (x22, x02) = x
        return [[self.add(0, x00, m[0]), x10], [x11, x21], [x22, self.add(2, x02, m[2])]]

    def rep_mir_sub(self, m, x):
        """SubOp::mir_rep_kernel (arith.rs:230-273): m - x."""
        (x00, x10), (x11, x21), (x22, x02) = x
        return [[self.sub(0, m[0], x00), self.neg(0, x10)], [self.neg(1, x11), self.neg(1, x21)],
                [self.neg(2, x22), self.sub(2, m[2], x02)]]

    def mir_fill(self, like_rep, value: int):
        """ShapeFill on a replicated tensor (misc.rs:136-170): a public constant on every player."""
        out = []
        for i in range(3):
            out.append(self.fill(i, self.shape(i, like_rep[i][0]), value, self.ty[like_rep[i][0]]))
        return out

    # ---- multiplication, dot, and (replicated/arith.rs:317-367, 436-508) ------------------------------------
    def _bilinear(self, op, x, y):
        v, shapes = [], []
        for i in range(3):
            (a0, a1), (b0, b1) = x[i], y[i]
            t = self.add(i, self.add(i, op(i, a0, b0), op(i, a0, b1)), op(i, a1, b0))
            v.append(t)
        for i in range(3):
            shapes.append(self.shape(i, v[i]))
        alphas = self.zero_share(shapes, self.ty[v[0]])
        z = [self.add(i, v[i], alphas[i]) for i in range(3)]
        return [[z[0], z[1]], [z[1], z[2]], [z[2], z[0]]]

    def rep_mul(self, x, y): return self._bilinear(self.mul, x, y)
    def rep_dot(self, x, y): return self._bilinear(self.dot, x, y)

    def rep_and(self, x, y):
        """AndOp::rep_kernel (arith.rs:495-508): the Z2 instance of mul (mul = and, add = xor, kernels/boolean.rs)."""
        v, shapes = [], []
        for i in range(3):
            (a0, a1), (b0, b1) = x[i], y[i]
            v.append(self.xor(i, self.xor(i, self.and_(i, a0, b0), self.and_(i, a0, b1)), self.and_(i, a1, b0)))
        for i in range(3):
            shapes.append(self.shape(i, v[i]))
        seeds = self.gen_seeds()
        z = []
        for i in range(3):
            r0 = self.sample_seeded(i, shapes[i], seeds[i][0], BIT_TY)
            r1 = self.sample_seeded(i, shapes[i], seeds[i][1], BIT_TY)
            z.append(self.xor(i, v[i], self.xor(i, r0, r1)))
        return [[z[0], z[1]], [z[1], z[2]], [z[2], z[0]]]

    # ---- truncation (replicated/fixedpoint.rs:80-103; additive/convert.rs:12-79, trunc.rs:22-169;
    #      replicated/convert.rs:392-497) -------------------------------------------------------------------
    def _adt_share_from_provider(self, v, key, shape_p, ty):
        """AdtShare at the provider (additive/trunc.rs:45-62): x0 = PRF(seed) and x1 = v - x0, both computed by the
        provider and moved to the additive players."""
        seed = self.derive_seed(2, key, self.sync_key())
        x0 = self.sample_seeded(2, shape_p, seed, ty)
        return [x0, self.sub(2, v, x0)]

    def rep_trunc_pr(self, amount: int, x):
        bits, ty = self.bits, self.ring_ty
        (x00, x10), (_x11, x21), _ = x
        a0, a1 = self.add(0, x00, x10), x21                               # rep_to_adt
        # gen_trunc_mask at the provider
        shape_p = self.shape(0, a0)
        r = self.sample(2, shape_p, ty)
        r_msb = self.shr(2, bits - 1, r)
        r_top = self.shr(2, amount + 1, self.shl(2, 1, r))
        key = self.key_gen(2)
        r_s = self._adt_share_from_provider(r, key, shape_p, ty)
        r_top_s = self._adt_share_from_provider(r_top, key, shape_p, ty)
        r_msb_s = self._adt_share_from_provider(r_msb, key, shape_p, ty)
        k = bits - 1
        ones = self.fill(0, shape_p, 1, ty)
        upshifter = self.shl(0, k - 1, ones)
        downshifter = self.shl(0, k - amount - 1, ones)
        xp0 = self.add(0, a0, upshifter)
        m0, m1 = self.add(0, xp0, r_s[0]), self.add(1, a1, r_s[1])
        c = self.add(0, m0, m1)                                           # adt reveal at player 0
        c_no_msb = self.shl(0, 1, c)
        c_top = self.shr(0, amount + 1, c_no_msb)
        c_msb = self.shr(0, bits - 1, c)
        t0, t1 = self.add(0, r_msb_s[0], c_msb), r_msb_s[1]
        # overflow = r_msb + c_msb - r_msb * c_msb - r_msb * c_msb: the product is evaluated twice, as written
        t0, t1 = self.sub(0, t0, self.mul(0, r_msb_s[0], c_msb)), self.sub(1, t1, self.mul(1, r_msb_s[1], c_msb))
        ov0, ov1 = self.sub(0, t0, self.mul(0, r_msb_s[0], c_msb)), self.sub(1, t1, self.mul(1, r_msb_s[1], c_msb))
        so0, so1 = self.shl(0, k - amount, ov0), self.shl(1, k - amount, ov1)
        y0, y1 = self.sub(0, c_top, r_top_s[0]), self.neg(1, r_top_s[1])
        yp0, yp1 = self.add(0, y0, so0), self.add(1, y1, so1)
        y_adt = [self.sub(0, yp0, downshifter), yp1]
        return self.adt_to_rep(y_adt)

    def adt_to_rep(self, x):
        x0, x1 = x
        ty = self.ty[x0]
        sk0, sk1 = self.sync_key(), self.sync_key()
        k = self.key_gen(2)
        seed1, seed2 = self.derive_seed(2, k, sk0), self.derive_seed(2, k, sk1)
        shape0, shape1 = self.shape(0, x0), self.shape(1, x1)
        y0 = self.sample_seeded(0, shape0, seed1, ty)
        y1 = self.sample_seeded(1, shape1, seed2, ty)
        y0_provider = self.sample_seeded(2, shape0, seed1, ty)
        y1_provider = self.sample_seeded(2, shape0, seed2, ty)
        c = self.add(0, self.sub(0, x0, y0), self.sub(1, x1, y1))        # opened at additive player 0, moved to 1
        return [[y0, c], [c, y1], [y1_provider, y0_provider]]

    # ---- arithmetic -> boolean (replicated/bits.rs:6-134, misc.rs:176-242) ----------------------------------
    def rep_split(self, x):
        (x00, x10), (_x11, x21), (x22, _x02) = x
        left = self.add(0, x00, x10)
        bsl = self.bit_decompose(0, left)
        x2_on_1, x2_on_2 = self.bit_decompose(1, x21), self.bit_decompose(2, x22)
        z0 = self.fill(0, self.shape(0, bsl), 0, BIT_TY)
        z1 = self.fill(1, self.shape(1, x2_on_1), 0, BIT_TY)
        z2 = self.fill(2, self.shape(2, x2_on_2), 0, BIT_TY)
        return self.rep_share(bsl, 0), [[z0, z0], [z1, x2_on_1], [x2_on_2, z2]]

    def binary_adder(self, x, y):
        n = self.bits
        g = self.rep_and(x, y)
        p_store = self.rep_xor(x, y)
        p = p_store
        for i in range(n.bit_length() - 1):
            p1 = self.rep_shl_dim(1 << i, n, p)
            g1 = self.rep_shl_dim(1 << i, n, g)
            g = self.rep_xor(g, self.rep_and(p, g1))
            p = self.rep_and(p, p1)
        return self.rep_xor(self.rep_shl_dim(1, n, g), p_store)

    def rep_bit_decompose(self, x):
        a, b = self.rep_split(x)
        return self.binary_adder(a, b)

    def rep_msb(self, x):
        return self.rep_index(self.bits - 1, self.rep_bit_decompose(x))

    # ---- Z2 -> ring (replicated/convert.rs:316-390; additive/dabit.rs:38-69) --------------------------------
    def rep_ring_inject(self, x, bit_idx: int = 0):
        ty = self.ring_ty
        (x00, x10), (_x11, x21), (x22, _x02) = x
        shape_p = self.shape(2, x22)
        b = self.sample(2, shape_p, BIT_TY)
        br = self.ring_inject(2, b, 0)
        key = self.key_gen(2)
        sk_b, sk_br = self.sync_key(), self.sync_key()
        seed_b, seed_br = self.derive_seed(2, key, sk_b), self.derive_seed(2, key, sk_br)
        b0 = self.sample_seeded(2, shape_p, seed_b, BIT_TY)
        b1 = self.xor(2, b, b0)
        br0 = self.sample_seeded(2, shape_p, seed_br, ty)
        br1 = self.sub(2, br, br0)
        xa0, xa1 = self.xor(0, x00, x10), x21                             # rep_to_adt over Z2
        c0, c1 = self.xor(0, xa0, b0), self.xor(1, xa1, b1)
        c_open = self.xor(0, c0, c1)
        c_ring = self.ring_inject(0, c_open, 0)
        t0, t1 = self.add(0, br0, c_ring), br1
        t0, t1 = self.sub(0, t0, self.mul(0, br0, c_ring)), self.sub(1, t1, self.mul(1, br1, c_ring))
        r0, r1 = self.sub(0, t0, self.mul(0, br0, c_ring)), self.sub(1, t1, self.mul(1, br1, c_ring))
        return self.adt_to_rep([self.shl(0, bit_idx, r0), self.shl(1, bit_idx, r1)])

    # ---- comparison, maximum (replicated/compare.rs:91-149, softmax.rs:10-53) --------------------------------
    def rep_less(self, x, y):
        return self.rep_msb(self.rep_sub(x, y))

    def rep_maximum(self, xs):
        n = len(xs)
        if n == 1:
            return xs[0]
        m1, m2 = self.rep_maximum(xs[: n // 2]), self.rep_maximum(xs[n // 2:])
        lesser_ring = self.rep_ring_inject(self.rep_less(m1, m2), 0)
        ones = self.mir_fill(lesser_ring, 1)
        a = self.rep_mul(lesser_ring, m2)
        b = self.rep_mul(self.rep_mir_sub(ones, lesser_ring), m1)
        return self.rep_add(a, b)

    # ---- whole computations ---------------------------------------------------------------------------------
    def fixed_dot_program(self, x_values, y_values, frac: int):
        """tutorials/dotprod.moose: x at player 0, y at player 1, fixed-point dot, revealed at player 2."""
        x = self.fixed_encode(0, self.constant(0, x_values), frac)
        y = self.fixed_encode(1, self.constant(1, y_values), frac)
        z = self.rep_trunc_pr(frac, self.rep_dot(self.rep_share(x, 0), self.rep_share(y, 1)))
        self.output(2, self.fixed_decode(2, self.rep_reveal(z, 2), frac), "output_0")
        return self.ops

    def mlp_program(self, n_layers: int, frac: int, reveal_to: int = 0):
        """pymoose MLPRegressor with ReLU (multilayer_perceptron_predictor.py:86-123): inputs `x` (player 0), mirrored
        `w{i}` / `b{i}` (one Input per player), zeros for maximum([zeros, z]) shared by player 1, logits revealed."""
        x = self.rep_share(self.fixed_encode(0, self.input(0, "x"), frac), 0)
        for l in range(n_layers):
            w_mir = [self.fixed_encode(i, self.input(i, f"w{l}"), frac) for i in range(3)]
            b_mir = [self.fixed_encode(i, self.input(i, f"b{l}"), frac) for i in range(3)]
            w = self.rep_share(w_mir[0], 0)                                    # mirrored -> replicated (mirrored/ops.rs:154-180)
            z = self.rep_add_mir(self.rep_trunc_pr(frac, self.rep_dot(x, w)), b_mir)
            if l < n_layers - 1:
                zeros = self.fill(1, self.shape(1, z[1][0]), 0, self.ring_ty)
                z = self.rep_maximum([self.rep_share(zeros, 1), z])
            x = z
        self.output(reveal_to, self.fixed_decode(reveal_to, self.rep_reveal(x, reveal_to), frac), "output_0")
        return self.ops

    def rep_index_axis(self, axis, index, x):
        return self._sharewise(lambda i, a: self.index_axis(i, a, axis, index), x)

    def rep_mux_bit(self, s_bits, x, y):
        """MuxOp::rep_bit_selector_kernel (replicated/control_flow.rs:27-44): ring_inject(s) * (x - y) + y."""
        s = self.rep_ring_inject(s_bits, 0)
        return self.rep_add(self.rep_mul(s, self.rep_sub(x, y)), y)

    def tree_regressor_program(self, trees, n_features: int, base_score: float, frac: int, learning_rate: float = 1.0,
                               reveal_to: int = 1):
        """pymoose TreeEnsembleRegressor (tree_ensemble.py:33-62, 83-114, 391-395) in the op order of
        moose_b200/predictors.py::TreeEnsemblePredictor: input `x` at player 0, shared by every index_axis; mirrored
        thresholds shared by player 0, leaf and base-score constants by player 2; add(base, add_n(scores)) revealed.
        Constants are one-element tensors (they broadcast against the [batch] columns like the reference's rank-0
        constants)."""
        xh = self.fixed_encode(0, self.input(0, "x"), frac)
        const = lambda h, v: self.fixed_encode(h, self.constant(h, [float(v)]), frac)

        def traverse(tree, node, feats):
            if tree.left[node] == 0 and tree.right[node] == 0:
                return None, const(2, learning_rate * tree.weights[node])
            thr = self.rep_share(const(0, tree.split_conditions[node]), 0)
            selector = self.rep_less(feats[tree.split_indices[node]], thr)
            lr, lh = traverse(tree, tree.left[node], feats)
            rr, rh = traverse(tree, tree.right[node], feats)
            xs = lr if lr is not None else self.rep_share(lh, 2)
            ys = rr if rr is not None else self.rep_share(rh, 2)
            return self.rep_mux_bit(selector, xs, ys), None

        total = None
        for tree in trees:
            feats = [self.rep_index_axis(1, i, self.rep_share(xh, 0)) for i in range(n_features)]
            score, leaf = traverse(tree, 0, feats)
            if score is None:
                raise NotImplementedError("degenerate tree (a single leaf)")
            total = score if total is None else self.rep_add(total, score)      # add_n
        y = self.rep_add(self.rep_share(const(2, base_score), 2), total)
        self.output(reveal_to, self.fixed_decode(reveal_to, self.rep_reveal(y, reveal_to), frac), "output_0")
        return self.ops


def prune_pass(ops: List[Operation]) -> List[Operation]:
    """Drops operations no Output / Save depends on (compilation/pruning.rs)."""
    by_name = {op.name: op for op in ops}
    live, stack = set(), [op.name for op in ops if op.kind in ("Output", "Save")]
    while stack:
        n = stack.pop()
        if n in live:
            continue
        live.add(n)
        stack.extend(i for i in by_name[n].inputs if i not in live)
    return [op for op in ops if op.name in live]


def networking_pass(ops: List[Operation]) -> List[Operation]:
    """One Send/Receive pair per (value, consuming host) for every edge that crosses hosts
    (compilation/networking.rs:75-118); rendezvous keys count up like the reference's."""
    by_name = {op.name: op for op in ops}
    out: List[Operation] = []
    moved: Dict[tuple, str] = {}
    n = 0
    for op in ops:
        dst = op.owners[0]
        new_inputs = []
        for i in op.inputs:
            src_op = by_name[i]
            src = src_op.owners[0]
            if src == dst:
                new_inputs.append(i)
                continue
            if (i, dst) not in moved:
                key = n.to_bytes(16, "little").hex()
                out.append(Operation(f"send_{n}", "Send", {"rendezvous_key": key, "receiver": f'"{dst}"'}, [src_op.output_type], False,
                                     "HostUnit", [i], "Host", [src]))
                out.append(Operation(f"receive_{n}", "Receive", {"rendezvous_key": key, "sender": f'"{src}"'}, [], False,
                                     src_op.output_type, [], "Host", [dst]))
                moved[(i, dst)] = f"receive_{n}"
                n += 1
            new_inputs.append(moved[(i, dst)])
        out.append(Operation(op.name, op.kind, op.attrs, op.input_types, op.variadic, op.output_type, new_inputs,
                             op.placement_kind, op.owners))
    return out
