"""GPU parity tests of the fused replicated protocols.  The CUDA library and the oracle are
given the same session id and entropy key; every replicated result is compared SHARE BY
SHARE (bit-exact), then revealed and compared with the plaintext expectation the
reference's own tests use (replicated/mod.rs, additive/trunc.rs)."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import BIT, RING64, RING128
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY, SESSION_ID, rand_ring, same, same_rep, same_ring

pytestmark = pytest.mark.gpu
DT = {64: RING64, 128: RING128}


def pair():
    return mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)


def signed(vals, bits):
    return [int(v) - (1 << bits) if int(v) >> (bits - 1) else int(v) for v in vals]


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("owner", [0, 1, 2, 3])
def test_share_reveal_parity(bits, owner):
    g, o = pair()
    rng = np.random.default_rng(bits + owner)
    for shape in [(5,), (3, 7), (), (130, 9)]:
        a = rand_ring(rng, bits, shape)
        gx = g.share(g.from_host(DT[bits], a), owner)
        ox = mo.rep_share(o, mo.Ring(bits, a), owner)
        assert same_rep(gx, ox)
        for recv in (0, 1, 2):
            r = g.reveal(gx, recv)
            assert same_ring(r, mo.rep_reveal(ox, recv))
            assert np.array_equal(r.to_host(), a)
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
def test_rep_linear_ops_parity(bits):
    g, o = pair()
    rng = np.random.default_rng(bits)
    a, b, p = rand_ring(rng, bits, (6, 5)), rand_ring(rng, bits, (6, 5)), rand_ring(rng, bits, (5,))
    gx, gy = g.share(g.from_host(DT[bits], a), 0), g.share(g.from_host(DT[bits], b), 1)
    ox, oy = mo.rep_share(o, mo.Ring(bits, a), 0), mo.rep_share(o, mo.Ring(bits, b), 1)
    gp, op = g.from_host(DT[bits], p), mo.Ring(bits, p)
    mir = mo.MirTensor([op, op, op])
    assert same_rep(g.rep_add(gx, gy), mo.rep_add(ox, oy))
    assert same_rep(g.rep_sub(gx, gy), mo.rep_sub(ox, oy))
    assert same_rep(g.rep_neg(gx), mo.rep_neg(ox))
    assert same_rep(g.rep_shl(7, gx), mo.rep_shl(7, ox))
    assert same_rep(g.rep_add_public(gx, gp), mo.rep_add_mir(ox, mir))
    assert same_rep(g.rep_sub_public(gx, gp), mo.rep_sub_mir(ox, mir))
    assert same_rep(g.rep_public_sub(gp, gx), mo.rep_mir_sub(mir, ox))
    assert same_rep(g.rep_mul_public(gx, gp), mo.rep_mul_mir(ox, mir))
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("shape", [(1,), (5,), (33, 3), (4, 3, 5), (1025,)])
def test_rep_mul_parity(bits, shape):
    """replicated/mod.rs:653-696 style wrap-around products, share-level parity."""
    g, o = pair()
    rng = np.random.default_rng(1)
    a, b = rand_ring(rng, bits, shape), rand_ring(rng, bits, shape)
    gx, gy = g.share(g.from_host(DT[bits], a), 0), g.share(g.from_host(DT[bits], b), 1)
    ox, oy = mo.rep_share(o, mo.Ring(bits, a), 0), mo.rep_share(o, mo.Ring(bits, b), 1)
    gz, oz = g.rep_mul(gx, gy), mo.rep_mul(o, ox, oy)
    assert same_rep(gz, oz)
    gz2, oz2 = g.rep_mul(gz, gx), mo.rep_mul(o, oz, ox)  # second mul: fresh sync keys, same setup
    assert same_rep(gz2, oz2)
    assert same_ring(g.reveal(gz), mo.mul(mo.Ring(bits, a), mo.Ring(bits, b)))
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("M,K,N", [(1, 3, 1), (4, 5, 6), (70, 33, 5), (130, 64, 66)])
def test_rep_dot_parity_imad(bits, M, K, N):
    g, o = pair()
    g.set_gemm_mode(1)
    rng = np.random.default_rng(M + K + N)
    a, b = rand_ring(rng, bits, (M, K)), rand_ring(rng, bits, (K, N))
    gx, gy = g.share(g.from_host(DT[bits], a), 0), g.share(g.from_host(DT[bits], b), 1)
    ox, oy = mo.rep_share(o, mo.Ring(bits, a), 0), mo.rep_share(o, mo.Ring(bits, b), 1)
    gz, oz = g.rep_dot(gx, gy), mo.rep_dot(o, ox, oy)
    assert same_rep(gz, oz)
    assert same_ring(g.reveal(gz), mo.dot(mo.Ring(bits, a), mo.Ring(bits, b)))
    g.close()


@pytest.mark.parametrize("bits,amount", [(64, 23), (64, 1), (64, 40), (128, 40), (128, 1), (128, 100)])
def test_trunc_pr_parity(bits, amount):
    """additive/trunc.rs:206-217,272-286 and replicated/mod.rs:847-859,965-980: |trunc - x>>amount| <= 1,
    plus bit-exact shares against the oracle under the shared entropy source."""
    g, o = pair()
    rng = np.random.default_rng(amount)
    k = bits - 1
    n = 1000
    vals = [int(v) for v in rng.integers(-(2 ** 62), 2 ** 62, n)]
    if bits == 128:
        vals = [v * int(rng.integers(1, 2 ** 62)) for v in vals]  # |x| < 2^124 <= 2^(k-2) holds for 125
        vals = [v >> 0 for v in vals]
    lim = 1 << (k - 2)
    vals = [max(-lim, min(lim - 1, v)) for v in vals] + [0, 1, -1, lim - 1, -lim]
    gx = g.share(g.from_ints(DT[bits], vals), 0)
    ox = mo.rep_share(o, mo.Ring.from_ints(bits, vals), 0)
    gz, oz = g.trunc_pr(amount, gx), mo.rep_trunc_pr(o, amount, ox)
    assert same_rep(gz, oz)
    got = signed(g.reveal(gz).to_ints(), bits)
    for v, t in zip(vals, got):
        assert abs(t - (v >> amount)) <= 1, (v, t, v >> amount)
    g.close()


def test_trunc_pr_rejects_large_amount():
    g, _ = pair()
    x = g.share(g.fill(RING64, 5, (3,)), 0)
    with pytest.raises(mb.MooseB200Error):
        g.trunc_pr(63, x)
    g.close()


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("shape", [(3,), (128,), (5, 7), (256,), (3, 128), (1024,)])
def test_bit_decompose_msb_parity(bits, shape, fused):
    """replicated/mod.rs:999-1032 (msb), :1100+ (bit_dec): share-level parity and plaintext check, for the
    fused per-level adder (n % 128 == 0; msb with carry pruning) and for the op-by-op composition."""
    g, o = pair()
    g.set_fused_bits(fused)
    rng = np.random.default_rng(9)
    a = rand_ring(rng, bits, shape)
    gx, ox = g.share(g.from_host(DT[bits], a), 0), mo.rep_share(o, mo.Ring(bits, a), 0)
    gb, ob = g.rep_bit_decompose(gx), mo.rep_bit_decompose(o, ox)
    assert same_rep(gb, ob)
    assert same(g.reveal(gb), mo.bit_decompose(mo.Ring(bits, a)))
    gm, om = g.msb(gx), mo.rep_msb(o, ox)
    assert same_rep(gm, om)
    assert same(g.reveal(gm), mo.bit_extract(bits - 1, mo.Ring(bits, a)))
    g.close()


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("bits", [64, 128])
def test_ring_inject_parity(bits, fused):
    """replicated/mod.rs ring_inject tests: b -> b << idx as a replicated ring value."""
    g, o = pair()
    g.set_fused_bits(fused)
    rng = np.random.default_rng(4)
    b = rng.integers(0, 2, (9, 5)).astype(np.uint8)
    gb, ob = g.share(g.from_bits(b), 0), mo.rep_share(o, mo.Bits.from_array(b), 0)
    assert same_rep(gb, ob)
    for idx in (0, 7, bits - 1):
        gr, orr = g.rep_ring_inject(DT[bits], idx, gb), mo.rep_ring_inject(o, bits, idx, ob)
        assert same_rep(gr, orr)
        assert [int(v) for v in g.reveal(gr).to_ints().reshape(-1)] == [int(v) << idx for v in b.reshape(-1)]
    g.close()


@pytest.mark.parametrize("bits,frac", [(64, 23), (128, 40)])
def test_relu_maximum_mux_parity(bits, frac):
    """fixedpoint/ops.rs:314-330 (relu), replicated/softmax.rs:10-53 (maximum), control_flow.rs (mux)."""
    g, o = pair()
    rng = np.random.default_rng(12)
    f = rng.uniform(-100, 100, (6, 7))
    vals = [int(v * 2 ** frac) for v in f.reshape(-1)]
    gx = g.share(g.from_ints(DT[bits], vals, f.shape), 0)
    ox = mo.rep_share(o, mo.Ring.from_ints(bits, vals, f.shape), 0)
    gr, orr = g.relu(gx), mo.rep_relu(o, ox)
    assert same_rep(gr, orr)
    assert signed(g.reveal(gr).to_ints().reshape(-1), bits) == [max(v, 0) for v in vals]
    # maximum([zeros, z]) as the MLP predictor builds it (zeros shared by bob)
    gz = g.share(g.fill(DT[bits], 0, f.shape), 1)
    oz = mo.rep_share(o, mo.fill(bits, 0, f.shape), 1)
    gm, om = g.maximum([gz, gx]), mo.rep_maximum(o, [oz, ox])
    assert same_rep(gm, om)
    assert signed(g.reveal(gm).to_ints().reshape(-1), bits) == [max(v, 0) for v in vals]
    w = [int(v) for v in rng.integers(-1000, 1000, f.size)]
    gw = g.share(g.from_ints(DT[bits], w, f.shape), 2)
    ow = mo.rep_share(o, mo.Ring.from_ints(bits, w, f.shape), 2)
    g3, o3 = g.maximum([gx, gw, gz]), mo.rep_maximum(o, [ox, ow, oz])
    assert same_rep(g3, o3)
    assert signed(g.reveal(g3).to_ints().reshape(-1), bits) == [max(a, b, 0) for a, b in zip(vals, w)]
    gl, ol = g.rep_less(gx, gw), mo.rep_less(o, ox, ow)
    assert same_rep(gl, ol)
    assert g.reveal(gl).bits().reshape(-1).tolist() == [int(a < b) for a, b in zip(vals, w)]
    g.close()


def test_fixed_dot_trunc_end_to_end():
    """tutorials/dotprod.moose: Fixed128(24,40) x=[[1,2,3]] . y=[[4],[5],[6]] = 32.0 (BASELINE config 1)."""
    g, o = pair()
    x = np.array([[1.0, 2.0, 3.0]])
    y = np.array([[4.0], [5.0], [6.0]])
    gx = g.share(g.fixedpoint_encode(RING128, 40, g.from_host(mb.FLOAT64, x)), 0)
    gy = g.share(g.fixedpoint_encode(RING128, 40, g.from_host(mb.FLOAT64, y)), 1)
    ox = mo.rep_share(o, mo.fixed_encode(128, x, 40), 0)
    oy = mo.rep_share(o, mo.fixed_encode(128, y, 40), 1)
    gz = g.trunc_pr(40, g.rep_dot(gx, gy))
    oz = mo.rep_trunc_pr(o, 40, mo.rep_dot(o, ox, oy))
    assert same_rep(gz, oz)
    out = g.fixedpoint_decode(40, g.reveal(gz)).to_host()
    assert abs(out[0, 0] - 32.0) < 1e-9
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("M,K,N", [(128, 128, 64), (256, 384, 320)])
def test_rep_dot_parity_tensor_pipe(bits, M, K, N):
    """Replicated dot through the two-term tensor-pipe GEMM: x_i.(y_i + y_{i+1}) + x_{i+1}.y_i."""
    g, o = pair()
    g.set_gemm_mode(2)
    rng = np.random.default_rng(M + K + N + 1)
    a, b = rand_ring(rng, bits, (M, K)), rand_ring(rng, bits, (K, N))
    gx, gy = g.share(g.from_host(DT[bits], a), 0), g.share(g.from_host(DT[bits], b), 1)
    ox, oy = mo.rep_share(o, mo.Ring(bits, a), 0), mo.rep_share(o, mo.Ring(bits, b), 1)
    gz, oz = g.rep_dot(gx, gy), mo.rep_dot(o, ox, oy)
    assert same_rep(gz, oz)
    assert same_ring(g.reveal(gz), mo.dot(mo.Ring(bits, a), mo.Ring(bits, b)))
    g.close()


@pytest.mark.parametrize("bits,frac", [(64, 23), (128, 40)])
def test_fixed_dot_and_fixed_mul_equal_the_two_call_sequence(bits, frac):
    """mb_rep_fixed_dot / _dot_concat / _fixed_mul = DotOp / MulOp followed by TruncPrOp (fixedpoint/ops.rs:761-928,
    548-672): the same entropy requests, the same result shares as the oracle's rep_dot / rep_mul + rep_trunc_pr, and
    the entropy source is left where the two-call sequence leaves it (the next op's shares agree too)."""
    g, o = pair()
    rng = np.random.default_rng(bits)
    for (M, K, N) in [(70, 33, 9), (256, 128, 64), (64, 513, 2)]:
        a, b = rng.uniform(-1, 1, (M, K)), rng.uniform(-1, 1, (K, N))
        gx = g.share(g.fixedpoint_encode(DT[bits], frac, g.from_host(mb.FLOAT64, a)), 0)
        gy = g.share(g.fixedpoint_encode(DT[bits], frac, g.from_host(mb.FLOAT64, b)), 1)
        ox, oy = mo.rep_share(o, mo.fixed_encode(bits, a, frac), 0), mo.rep_share(o, mo.fixed_encode(bits, b, frac), 1)
        gz = g.fixed_dot(frac, gx, gy)
        oz = mo.rep_trunc_pr(o, frac, mo.rep_dot(o, ox, oy))
        assert same_rep(gz, oz)
        gz2 = g.fixed_dot_concat(frac, [gx, gx], g.rep_concat(0, [gy, gy]))
        oz2 = mo.rep_trunc_pr(o, frac, mo.rep_dot(o, mo.rep_concat(1, [ox, ox]), mo.rep_concat(0, [oy, oy])))
        assert same_rep(gz2, oz2)
        gm = g.fixed_mul(frac, gz, gz2)
        om = mo.rep_trunc_pr(o, frac, mo.rep_mul(o, oz, oz2))
        assert same_rep(gm, om)
        assert np.abs(g.fixedpoint_decode(frac, g.reveal(gz)).to_host() - a @ b).max() < 1e-3
    # broadcasting operands fall back to the two-call sequence, still the same shares
    gs = g.share(g.fixedpoint_encode(DT[bits], frac, g.from_host(mb.FLOAT64, np.array([[0.5]]))), 2)
    os_ = mo.rep_share(o, mo.fixed_encode(bits, np.array([[0.5]]), frac), 2)
    assert same_rep(g.fixed_mul(frac, gz, gs), mo.rep_trunc_pr(o, frac, mo.rep_mul(o, oz, os_)))
    assert same_rep(g.rep_mul(gz, gz), mo.rep_mul(o, oz, oz))       # entropy positions still agree afterwards
    g.close()


@pytest.mark.parametrize("bits,frac,N", [(128, 40, 2), (64, 23, 4), (64, 23, 3), (128, 40, 5)])
@pytest.mark.parametrize("owner", [0, 1, 2])
def test_deferred_share_expanded_inside_the_skinny_dot(bits, frac, N, owner):
    """mb_rep_share_deferred + mb_rep_(fixed_)dot_concat: the owner's PRF share is expanded inside the dot kernel (or,
    where no kernel takes it -- N = 5 on Ring128 here -- written out first); the result shares are those of ShareOp +
    ConcatOp + DotOp (+ TruncPrOp) on the oracle, and accessing a share of the deferred handle materialises exactly
    what ShareOp writes."""
    g, o = pair()
    rng = np.random.default_rng(bits + N + owner)
    M, K = 96, 34
    a, b = rng.uniform(-1, 1, (M, K)), rng.uniform(-1, 1, (K + 1, N))
    ones = np.ones((M, 1))
    ga, gb, g1 = (g.fixedpoint_encode(DT[bits], frac, g.from_host(mb.FLOAT64, v)) for v in (a, b, ones))
    oa, ob, o1 = (mo.fixed_encode(bits, v, frac) for v in (a, b, ones))
    gbias, obias = g.share(g1, 1), mo.rep_share(o, o1, 1)
    gx = g.share(ga, owner, deferred=True)
    ox = mo.rep_share(o, oa, owner)
    gw, ow = g.share(gb, 0), mo.rep_share(o, ob, 0)
    with pytest.raises(mb.MooseB200Error):
        g.rep_add(gx, gx)                       # a deferred handle is only accepted by its consumers
    gz = g.fixed_dot_concat(frac, [gbias, gx], gw)
    oz = mo.rep_trunc_pr(o, frac, mo.rep_dot(o, mo.rep_concat(1, [obias, ox]), ow))
    assert same_rep(gz, oz)
    gz2 = g.rep_dot_concat([gbias, gx], gw)     # with the zero share
    oz2 = mo.rep_dot(o, mo.rep_concat(1, [obias, ox]), ow)
    assert same_rep(gz2, oz2)
    assert same_rep(gx, ox)                     # materialised on access: exactly ShareOp's tensors
    assert same_rep(g.rep_add(gx, gx), mo.rep_add(ox, ox))
    g.close()


@pytest.mark.parametrize("N", [2, 10, 16])
def test_rep_dot_tall_left_operand_narrow_right_ring64(N):
    """[M, K] x [K, N] with M >= 256 rows per SM and an even N <= 16 on Ring64 takes the one-lane-per-row kernel
    (csrc/gemm_imad.cu, rep_dot_rowlane_kernel: the last layer of the MLP): share-level parity, ragged M."""
    g, o = pair()
    rng = np.random.default_rng(N)
    M, K = 256 * 148 + 37, 24
    a, b = rand_ring(rng, 64, (M, K)), rand_ring(rng, 64, (K, N))
    gx, gy = g.share(g.from_host(RING64, a), 0), g.share(g.from_host(RING64, b), 2)
    ox, oy = mo.rep_share(o, mo.Ring(64, a), 0), mo.rep_share(o, mo.Ring(64, b), 2)
    assert same_rep(g.rep_dot(gx, gy), mo.rep_dot(o, ox, oy))
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
def test_mux_with_ring_and_bit_selectors(bits):
    """replicated/control_flow.rs:53-123 (test_mux, test_bit_selector_mux): s*(x-y)+y with the selector as a ring
    value and as replicated bits (ring_inject first), share-level parity and the reference's decoded answers."""
    g, o = pair()
    enc = lambda v: [int(t * 2 ** 24) for t in v]
    for sel, xs, ys, want in [([1, 0, 1, 0], [1.0, 2.0, 3.0, 4.0], [4.0, 5.0, 6.0, -1.0], [1.0, 5.0, 3.0, -1.0]),
                              ([1, 1, 0], [1.0, 2.0, 3.0], [4.0, 5.0, 6.0], [1.0, 2.0, 6.0])]:
        n = len(sel)
        gx, ox = g.share(g.from_ints(DT[bits], enc(xs), (n,)), 0), mo.rep_share(o, mo.Ring.from_ints(bits, enc(xs)), 0)
        gy, oy = g.share(g.from_ints(DT[bits], enc(ys), (n,)), 1), mo.rep_share(o, mo.Ring.from_ints(bits, enc(ys)), 1)
        b = np.array(sel, dtype=np.uint8)
        gs, os_ = g.share(g.from_bits(b), 0), mo.rep_share(o, mo.Bits.from_array(b), 0)
        gm, om = g.mux(gs, gx, gy), mo.rep_mux_bit_selector(o, bits, os_, ox, oy)
        assert same_rep(gm, om)
        assert [v / 2 ** 24 for v in signed(g.reveal(gm).to_ints().reshape(-1), bits)] == want
        gr, orr = g.share(g.from_ints(DT[bits], sel, (n,)), 2), mo.rep_share(o, mo.Ring.from_ints(bits, sel), 2)
        gm, om = g.mux(gr, gx, gy), mo.rep_mux(o, orr, ox, oy)
        assert same_rep(gm, om)
        assert [v / 2 ** 24 for v in signed(g.reveal(gm).to_ints().reshape(-1), bits)] == want
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("shape", [(3,), (128,), (5, 7)])
def test_bit_compose_inverts_bit_decompose(bits, shape):
    """replicated/mod.rs:1107-1139 (rep_bit_compose tests): compose(decompose(x)) == x, every share equal to the
    oracle's fold of k ring_injects."""
    g, o = pair()
    rng = np.random.default_rng(21)
    a = rand_ring(rng, bits, shape)
    if shape == (3,):
        a = mo.Ring.from_ints(bits, [1073741823, 0, 6]).data
    gx, ox = g.share(g.from_host(DT[bits], a), 0), mo.rep_share(o, mo.Ring(bits, a), 0)
    gb, ob = g.rep_bit_decompose(gx), mo.rep_bit_decompose(o, ox)
    gc, oc = g.bit_compose(DT[bits], gb), mo.rep_bit_compose(o, bits, ob)
    assert same_rep(gc, oc)
    assert np.array_equal(g.reveal(gc).to_host(), a)
    with pytest.raises(mb.MooseB200Error):
        g.bit_compose(DT[bits], gx)          # not bit planes
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("shape", [(0,), (0, 5), (3, 0)])
def test_empty_tensors_through_the_replicated_layer(bits, shape):
    """Zero-element tensors (ndarray handles them; the reference's kernels are shape-generic): every protocol returns
    the empty shape, draws the same entropy requests as on the oracle -- checked by a non-empty multiplication after
    them, share by share -- and a dot with an empty contraction axis is the zero matrix plus the zero share."""
    g, o = pair()
    a = np.zeros(shape + ((2,) if bits == 128 else ()), dtype=np.uint64)
    gx, ox = g.share(g.from_host(DT[bits], a), 0), mo.rep_share(o, mo.Ring(bits, a), 0)
    gy, oy = g.share(g.from_host(DT[bits], a), 1), mo.rep_share(o, mo.Ring(bits, a), 1)
    steps = [
        (lambda: g.rep_add(gx, gy), lambda: mo.rep_add(ox, oy)),
        (lambda: g.rep_mul(gx, gy), lambda: mo.rep_mul(o, ox, oy)),
        (lambda: g.trunc_pr(5, gx), lambda: mo.rep_trunc_pr(o, 5, ox)),
        (lambda: g.msb(gx), lambda: mo.rep_msb(o, ox)),
        (lambda: g.rep_bit_decompose(gx), lambda: mo.rep_bit_decompose(o, ox)),
        (lambda: g.relu(gx), lambda: mo.rep_relu(o, ox)),
        (lambda: g.maximum([gx, gy]), lambda: mo.rep_maximum(o, [ox, oy])),
        (lambda: g.rep_less(gx, gy), lambda: mo.rep_less(o, ox, oy)),
    ]
    for gf, of in steps:
        gr, orr = gf(), of()
        assert same_rep(gr, orr)
    assert g.reveal(g.relu(gx)).shape == shape and mo.rep_reveal(mo.rep_relu(o, ox)).shape == shape
    if len(shape) == 2:
        rng = np.random.default_rng(3)
        other = (shape[1], 2)
        b = rand_ring(rng, bits, other)
        gb, ob = g.share(g.from_host(DT[bits], b), 2), mo.rep_share(o, mo.Ring(bits, b), 2)
        gd, od = g.rep_dot(gx, gb), mo.rep_dot(o, ox, ob)
        assert same_rep(gd, od)
        r = g.reveal(gd)
        assert r.shape == (shape[0], 2) and not r.to_host().any()
    # entropy still in step: a non-empty product after the empty calls
    rng = np.random.default_rng(4)
    p, q = rand_ring(rng, bits, (9,)), rand_ring(rng, bits, (9,))
    gp, op = g.share(g.from_host(DT[bits], p), 0), mo.rep_share(o, mo.Ring(bits, p), 0)
    gq, oq = g.share(g.from_host(DT[bits], q), 1), mo.rep_share(o, mo.Ring(bits, q), 1)
    assert same_rep(g.rep_mul(gp, gq), mo.rep_mul(o, op, oq))
    g.close()
