"""oracle/moose_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of the reference's replicated / additive / mirrored protocol
layer, op for op, on top of the C restatement of the host kernels
(oracle/host_kernels.c).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module; nothing under
moose_b200/ does.

Every routine cites the reference code it follows (paths relative to
/root/reference/moose/src/).  The reference keeps six share tensors per
replicated value, `shares = [[x00, x10], [x11, x21], [x22, x02]]`
(replicated/mod.rs:75-77): party i holds (x_i, x_{i+1}).  This oracle keeps
that layout verbatim so the op order -- and therefore the order in which
entropy is consumed -- is the reference's.

Determinism contract.  The reference draws PRF
keys, unseeded Sample seeds and sync keys from OS entropy, so its shares are
not reproducible.  Here all three come from one explicit `EntropySource`
(AES-128-CTR keyed by a 16-byte test key, one 16-byte block per request),
consumed in the reference's program order.  The CUDA library implements the
same contract, which is what makes share-level bit-exact parity testable.

Parity pinning: AES-CTR u64/u128 streams, DeriveSeed and SessionId hashing are
pinned by the reference's golden vectors (tests/test_oracle_kat.py).  The
get_bit() policy is pinned by 5 bits only -> "parity partially pinned" for raw
bit-share streams; revealed values do not depend on it.  The protocol layer is
pinned end to end by the reference's compiled graphs (tests/test_textual.py)
and by the known answers of its predictor tests on its own ONNX fixtures
(tests/test_tree_ensembles.py, test_linear_golden.py, test_mlp_golden.py).
"""
from __future__ import annotations

import ctypes
import math
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

try:  # blake3 is an independent third-party implementation (python wheel)
    import blake3 as _blake3
except Exception:  # pragma: no cover
    _blake3 = None

from . import build as _build

_lib = ctypes.CDLL(_build.build())

_u8p = ctypes.POINTER(ctypes.c_uint8)
_vp = ctypes.c_void_p
_u64 = ctypes.c_uint64


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_vp)


def _sig(name, *args):
    f = getattr(_lib, name)
    f.argtypes = list(args)
    f.restype = None
    return f


_aes_ctr = _sig("orc_aes_ctr", _vp, _u64, _u64, _vp)
_aes_block = _sig("orc_aes_block", _vp, _vp, _vp)
_sample_ring64 = _sig("orc_sample_ring64", _vp, _u64, _vp)
_sample_ring128 = _sig("orc_sample_ring128", _vp, _u64, _vp)
_sample_bits = _sig("orc_sample_bits", _vp, _u64, ctypes.c_int, _vp)
_sample_ring64_bits = _sig("orc_sample_ring64_bits", _vp, _u64, ctypes.c_int, _vp)
_sample_ring128_bits = _sig("orc_sample_ring128_bits", _vp, _u64, ctypes.c_int, _vp)
_force_portable = _sig("orc_force_portable_aes", ctypes.c_int)
_lib.orc_max_threads.restype = ctypes.c_int
for _w in ("64", "128"):
    for _op in ("add", "sub", "mul", "and"):
        _sig(f"orc_{_op}{_w}", _vp, _vp, _vp, _u64)
    _sig(f"orc_neg{_w}", _vp, _vp, _u64)
    _sig(f"orc_shl{_w}", _vp, _vp, _u64, _u64)
    _sig(f"orc_shr{_w}", _vp, _vp, _u64, _u64)
    _sig(f"orc_dot{_w}", _vp, _vp, _vp, _u64, _u64, _u64, ctypes.c_int)
    _sig(f"orc_sum_axis{_w}", _vp, _vp, _u64, _u64, _u64)
    _sig(f"orc_bit_decompose{_w}", _vp, _u64, _vp)
    _sig(f"orc_bit_extract{_w}", _vp, _u64, _u64, _vp)
    _sig(f"orc_ring_inject{_w}", _vp, _u64, _u64, _vp)
    _sig(f"orc_fixed_encode{_w}", _vp, _u64, ctypes.c_double, _vp)
    _sig(f"orc_fixed_decode{_w}", _vp, _u64, ctypes.c_double, _vp)
for _op in ("xor", "and", "or"):
    _sig(f"orc_bit_{_op}", _vp, _vp, _vp, _u64)
_sig("orc_bit_not", _vp, _vp, _u64)
_sig("orc_shl_dim", _vp, _u64, _u64, _u64, _vp)
_sig("orc_bit_copy_range", _vp, _u64, _u64, _vp)
_sig("orc_bit_pack", _vp, _u64, _vp)
_sig("orc_bit_unpack", _vp, _u64, _vp)

DOT_THREADS = 1  # reference host kernels are single threaded


def max_threads() -> int:
    return int(_lib.orc_max_threads())


# ---------------------------------------------------------------------------
# primitives: AES-CTR stream, BLAKE3 seed derivation, session id
# ---------------------------------------------------------------------------

def aes_ctr(seed: bytes, start_block: int, nblocks: int) -> bytes:
    """aes-prng stream: block i = AES-128_seed(LE128(i))."""
    out = np.empty(16 * nblocks, dtype=np.uint8)
    s = np.frombuffer(bytes(seed), dtype=np.uint8).copy()
    _aes_ctr(_ptr(s), start_block, nblocks, _ptr(out))
    return out.tobytes()


def aes_block(key: bytes, block: bytes) -> bytes:
    out = np.empty(16, dtype=np.uint8)
    k = np.frombuffer(bytes(key), dtype=np.uint8).copy()
    b = np.frombuffer(bytes(block), dtype=np.uint8).copy()
    _aes_block(_ptr(k), _ptr(b), _ptr(out))
    return out.tobytes()


def session_id_bytes(logical: str) -> bytes:
    """SessionId::try_from(&str) (computation.rs:108-124): blake3(s).xof(16)."""
    return _blake3.blake3(logical.encode()).digest(length=16)


def sync_key_from(values: Sequence[int]) -> bytes:
    """SyncKey::try_from(&[u8]) (host/prim.rs:98-111): zero padded to 16."""
    b = bytes(values)
    assert len(b) <= 16
    return b + bytes(16 - len(b))


def derive_seed(sid: bytes, sync_key: bytes, prf_key: bytes) -> bytes:
    """DeriveSeedOp::kernel (host/prim.rs:123-149)."""
    dk = _blake3.blake3(prf_key, derive_key_context="Derive Seed").digest()
    h = _blake3.blake3(key=dk)
    h.update(sid)
    h.update(sync_key)
    return h.digest(length=16)


class EntropySource:
    """Explicit stand-in for the reference's OS entropy (determinism contract).

    Request i returns AES-128_key(LE128(i)).  Consumed by PrfKeyGen
    (host/prim.rs:118), unseeded Sample (host/ops.rs:1890,1917,1951: the 16
    bytes are that op's seed) and SyncKey::random() (host/prim.rs:73-79).
    """

    def __init__(self, key: bytes):
        assert len(key) == 16
        self.key = bytes(key)
        self.counter = 0

    def next16(self) -> bytes:
        out = aes_ctr(self.key, self.counter, 1)
        self.counter += 1
        return out


# ---------------------------------------------------------------------------
# host tensors
# ---------------------------------------------------------------------------

class Ring:
    """HostRingTensor<u64|u128> (host/mod.rs:388). 128-bit data is uint64[...,2] = [lo, hi]."""

    def __init__(self, bits: int, data: np.ndarray):
        assert bits in (64, 128)
        self.bits = bits
        shp = np.shape(data)
        data = np.ascontiguousarray(data, dtype=np.uint64).reshape(shp)  # keep 0-d shapes
        if bits == 128:
            assert data.shape[-1] == 2
        self.data = data

    @property
    def shape(self) -> Tuple[int, ...]:
        return tuple(self.data.shape[:-1]) if self.bits == 128 else tuple(self.data.shape)

    @property
    def size(self) -> int:
        return int(np.prod(self.shape, dtype=np.int64)) if len(self.shape) else 1

    @staticmethod
    def from_ints(bits: int, values, shape=None) -> "Ring":
        arr = np.array(values, dtype=object)
        if shape is None:
            shape = arr.shape
        flat = [int(v) % (1 << bits) for v in arr.reshape(-1)]
        if bits == 64:
            return Ring(64, np.array(flat, dtype=np.uint64).reshape(shape))
        lo = np.array([v & ((1 << 64) - 1) for v in flat], dtype=np.uint64)
        hi = np.array([v >> 64 for v in flat], dtype=np.uint64)
        return Ring(128, np.stack([lo, hi], axis=-1).reshape(tuple(shape) + (2,)))

    def to_ints(self) -> np.ndarray:
        if self.bits == 64:
            return np.array([int(v) for v in self.data.reshape(-1)], dtype=object).reshape(self.shape)
        d = self.data.reshape(-1, 2)
        return np.array([int(lo) | (int(hi) << 64) for lo, hi in d], dtype=object).reshape(self.shape)

    def reshape(self, shape) -> "Ring":
        shape = tuple(shape)
        return Ring(self.bits, self.data.reshape(shape + ((2,) if self.bits == 128 else ())))

    def __eq__(self, other):
        return isinstance(other, Ring) and self.bits == other.bits and self.shape == other.shape \
            and bool(np.array_equal(self.data, other.data))

    def tobytes(self) -> bytes:
        return self.data.tobytes()


class Bits:
    """HostBitTensor (host/mod.rs:205; bitarray.rs:10-13): packed Lsb0 bytes + shape."""

    def __init__(self, shape: Sequence[int], packed: np.ndarray):
        self.shape = tuple(int(s) for s in shape)
        n = self.size
        packed = np.ascontiguousarray(packed, dtype=np.uint8)
        assert packed.size == (n + 7) // 8, (packed.size, n)
        self.packed = packed

    @property
    def size(self) -> int:
        return int(np.prod(self.shape, dtype=np.int64)) if len(self.shape) else 1

    @staticmethod
    def from_array(values) -> "Bits":
        a = np.asarray(values, dtype=np.uint8)
        flat = np.ascontiguousarray(a.reshape(-1) & 1)
        out = np.zeros((flat.size + 7) // 8, dtype=np.uint8)
        _lib.orc_bit_pack(_ptr(flat), flat.size, _ptr(out))
        return Bits(a.shape, out)

    def to_array(self) -> np.ndarray:
        out = np.zeros(max(self.size, 1), dtype=np.uint8)
        _lib.orc_bit_unpack(_ptr(self.packed), self.size, _ptr(out))
        return out[: self.size].reshape(self.shape)

    def __eq__(self, other):
        return isinstance(other, Bits) and self.shape == other.shape \
            and bool(np.array_equal(self.packed, other.packed))


def _w(x: Ring) -> str:
    return str(x.bits)


def _bcast(x: Ring, y: Ring) -> Tuple[Ring, Ring, Tuple[int, ...]]:
    """ndarray co-broadcasting in `x.0 + y.0` (host/mod.rs:813-829)."""
    assert x.bits == y.bits
    shape = np.broadcast_shapes(x.shape, y.shape)
    if x.shape != shape:
        x = broadcast(x, shape)
    if y.shape != shape:
        y = broadcast(y, shape)
    return x, y, shape


def broadcast(x: Ring, shape) -> Ring:
    """BroadcastOp::host_ring_kernel (host/ops.rs:1668-1681)."""
    shape = tuple(shape)
    tail = (2,) if x.bits == 128 else ()
    return Ring(x.bits, np.ascontiguousarray(np.broadcast_to(x.data, shape + tail)).reshape(shape + tail))


def _ew2(op: str, x, y):
    if isinstance(x, Bits):
        # Z2: add = sub = xor, mul = and (kernels/boolean.rs:18-76)
        bop = {"add": "xor", "sub": "xor", "mul": "and"}[op]
        return bit_op(bop, x, y)
    x, y, shape = _bcast(x, y)
    out = np.empty_like(x.data)
    getattr(_lib, f"orc_{op}{_w(x)}")(_ptr(x.data), _ptr(y.data), _ptr(out), x.size)
    return Ring(x.bits, out)


def add(x, y):
    """AddOp::ring_kernel (host/ops.rs:1709-1722)."""
    return _ew2("add", x, y)


def sub(x, y):
    """SubOp::ring_kernel (host/ops.rs:1724-1737)."""
    return _ew2("sub", x, y)


def mul(x, y):
    """MulOp::ring_kernel (host/ops.rs:299-310)."""
    return _ew2("mul", x, y)


def neg(x):
    """NegOp::ring_kernel (host/ops.rs:1739-1752) / NegOp::bit_kernel (:1555-1565)."""
    if isinstance(x, Bits):
        out = np.empty_like(x.packed)
        _lib.orc_bit_not(_ptr(x.packed), _ptr(out), x.size)
        return Bits(x.shape, out)
    out = np.empty_like(x.data)
    getattr(_lib, f"orc_neg{_w(x)}")(_ptr(x.data), _ptr(out), x.size)
    return Ring(x.bits, out)


def shl(amount: int, x: Ring) -> Ring:
    """ShlOp::ring_kernel (host/ops.rs:1853-1866); amount masked mod k (Wrapping)."""
    out = np.empty_like(x.data)
    getattr(_lib, f"orc_shl{_w(x)}")(_ptr(x.data), _ptr(out), x.size, amount)
    return Ring(x.bits, out)


def shr(amount: int, x: Ring) -> Ring:
    """ShrOp::ring_kernel (host/ops.rs:1868-1881)."""
    out = np.empty_like(x.data)
    getattr(_lib, f"orc_shr{_w(x)}")(_ptr(x.data), _ptr(out), x.size, amount)
    return Ring(x.bits, out)


def dot(x: Ring, y: Ring) -> Ring:
    """HostRingTensor::dot / DotOp::ring_kernel (host/ops.rs:1754-1851)."""
    assert x.bits == y.bits
    rx, ry = len(x.shape), len(y.shape)
    if rx not in (1, 2) or ry not in (1, 2):
        raise ValueError(f"Dot<HostRingTensor> cannot handle ranks {rx},{ry}")
    M = x.shape[0] if rx == 2 else 1
    K = x.shape[-1]
    N = y.shape[1] if ry == 2 else 1
    if y.shape[0] != K:
        raise ValueError("dot: inner dimensions differ")
    tail = (2,) if x.bits == 128 else ()
    out = np.empty((M, N) + tail, dtype=np.uint64)
    getattr(_lib, f"orc_dot{_w(x)}")(_ptr(x.data), _ptr(y.data), _ptr(out), M, K, N, DOT_THREADS)
    if rx == 2 and ry == 2:
        oshape = (M, N)
    elif rx == 2:
        oshape = (M,)
    elif ry == 2:
        oshape = (N,)
    else:
        oshape = ()
    return Ring(x.bits, out.reshape(oshape + tail))


def add_n(xs: Sequence[Ring]) -> Ring:
    """AddNOp::host_kernel (host/ops.rs:1084-1103)."""
    acc = xs[0]
    for v in xs[1:]:
        acc = add(acc, v)
    return acc


def sum_axis(x: Ring, axis: Optional[int]) -> Ring:
    """SumOp::host_ring_kernel / HostRingTensor::sum (host/ops.rs:1067-1081; host/mod.rs:476-487)."""
    shape = x.shape
    tail = (2,) if x.bits == 128 else ()
    if axis is None:
        outer, ln, inner, oshape = 1, x.size, 1, ()
    else:
        outer = int(np.prod(shape[:axis], dtype=np.int64))
        ln = shape[axis]
        inner = int(np.prod(shape[axis + 1:], dtype=np.int64))
        oshape = shape[:axis] + shape[axis + 1:]
    out = np.empty((max(outer * inner, 1),) + tail, dtype=np.uint64)
    getattr(_lib, f"orc_sum_axis{_w(x)}")(_ptr(x.data), _ptr(out), outer, ln, inner)
    return Ring(x.bits, out.reshape(tuple(oshape) + tail))


def fill(bits: int, value: int, shape) -> Ring:
    """FillOp::host_ring64/128_kernel (host/ops.rs:1635-1655)."""
    shape = tuple(shape)
    v = int(value) % (1 << bits)
    if bits == 64:
        return Ring(64, np.full(shape, v, dtype=np.uint64))
    out = np.empty(shape + (2,), dtype=np.uint64)
    out[..., 0] = v & ((1 << 64) - 1)
    out[..., 1] = v >> 64
    return Ring(128, out)


def fill_bits(value: int, shape) -> Bits:
    """FillOp::host_bit_kernel (host/ops.rs:1623-1633)."""
    shape = tuple(shape)
    n = int(np.prod(shape, dtype=np.int64)) if len(shape) else 1
    return Bits.from_array(np.full(shape, value & 1, dtype=np.uint8)) if n else Bits(shape, np.zeros(0, np.uint8))


def zeros_like(x):
    return fill_bits(0, x.shape) if isinstance(x, Bits) else fill(x.bits, 0, x.shape)


def bit_op(op: str, x: Bits, y: Bits) -> Bits:
    """XorOp/AndOp/OrOp::host_kernel (host/ops.rs:1507-1553)."""
    assert x.shape == y.shape, (x.shape, y.shape)
    out = np.empty_like(x.packed)
    getattr(_lib, f"orc_bit_{op}")(_ptr(x.packed), _ptr(y.packed), _ptr(out), x.size)
    return Bits(x.shape, out)


def ring_and(x: Ring, y: Ring) -> Ring:
    """AndOp on ring tensors (kernels/boolean.rs): bitwise and."""
    x, y, _ = _bcast(x, y)
    out = np.empty_like(x.data)
    getattr(_lib, f"orc_and{_w(x)}")(_ptr(x.data), _ptr(y.data), _ptr(out), x.size)
    return Ring(x.bits, out)


def bit_decompose(x: Ring) -> Bits:
    """BitDecomposeOp::host_bit64/128_kernel (host/ops.rs:896-936): [k, shape...]."""
    n = x.size
    out = np.zeros((x.bits * n + 7) // 8, dtype=np.uint8)
    getattr(_lib, f"orc_bit_decompose{_w(x)}")(_ptr(x.data), n, _ptr(out))
    return Bits((x.bits,) + x.shape, out)


def bit_extract(bit_idx: int, x: Ring) -> Bits:
    """BitExtractOp::kernel64/128 (host/ops.rs:1567-1595)."""
    n = x.size
    out = np.zeros((n + 7) // 8, dtype=np.uint8)
    getattr(_lib, f"orc_bit_extract{_w(x)}")(_ptr(x.data), n, bit_idx, _ptr(out))
    return Bits(x.shape, out)


def ring_inject(bits: int, bit_idx: int, x: Bits) -> Ring:
    """RingInjectOp::host_kernel (host/ops.rs:1597-1621)."""
    n = x.size
    tail = (2,) if bits == 128 else ()
    out = np.empty((max(n, 1),) + tail, dtype=np.uint64)
    getattr(_lib, f"orc_ring_inject{bits}")(_ptr(x.packed), n, bit_idx, _ptr(out))
    return Ring(bits, out[:n].reshape(x.shape + tail))


def shl_dim(amount: int, bit_length: int, x: Bits) -> Bits:
    """ShlDimOp::host_bit_kernel (host/ops.rs:835-853)."""
    height = x.size // x.shape[0] if x.shape[0] else 0
    out = np.zeros_like(x.packed)
    _lib.orc_shl_dim(_ptr(x.packed), height, bit_length, amount, _ptr(out))
    return Bits(x.shape, out)


def index_axis0(index: int, x):
    """IndexAxisOp with axis 0 (host/ops.rs:740-783; bitarray.rs:85-108)."""
    if isinstance(x, Bits):
        height = x.size // x.shape[0]
        out = np.zeros((height + 7) // 8, dtype=np.uint8)
        _lib.orc_bit_copy_range(_ptr(x.packed), index * height, height, _ptr(out))
        return Bits(x.shape[1:], out)
    return Ring(x.bits, x.data[index].copy())


def fixed_encode(bits: int, x: np.ndarray, frac: int) -> Ring:
    """RingFixedpointEncodeOp (host/ops.rs:1377-1389; host/fixedpoint.rs:10-38)."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    tail = (2,) if bits == 128 else ()
    out = np.empty(x.shape + tail, dtype=np.uint64)
    getattr(_lib, f"orc_fixed_encode{bits}")(_ptr(x), x.size, float(2 ** frac), _ptr(out))
    return Ring(bits, out)


def fixed_decode(x: Ring, frac: int) -> np.ndarray:
    """RingFixedpointDecodeOp (host/ops.rs:1403-1417)."""
    out = np.empty(x.shape, dtype=np.float64)
    getattr(_lib, f"orc_fixed_decode{_w(x)}")(_ptr(x.data), x.size, float(2 ** frac), _ptr(out))
    return out


# -- sampling ----------------------------------------------------------------

BIT_POLICY_STREAM_LSB0 = 0
BIT_POLICY_U16_PER_BIT = 1


def sample_uniform_seeded(bits: int, shape, seed: bytes) -> Ring:
    """SampleSeededOp::ring64/128_kernel, max_value=None (host/ops.rs:1959-2017)."""
    shape = tuple(shape)
    n = int(np.prod(shape, dtype=np.int64)) if len(shape) else 1
    tail = (2,) if bits == 128 else ()
    out = np.empty((max(n, 1),) + tail, dtype=np.uint64)
    s = np.frombuffer(bytes(seed), dtype=np.uint8).copy()
    (_sample_ring64 if bits == 64 else _sample_ring128)(_ptr(s), n, _ptr(out))
    return Ring(bits, out[:n].reshape(shape + tail))


def sample_bits_seeded_ring(bits: int, shape, seed: bytes, policy: int = 0) -> Ring:
    """SampleSeededOp ring kernels with max_value=Some(1) (host/ops.rs:1971-1974,2001-2004)."""
    shape = tuple(shape)
    n = int(np.prod(shape, dtype=np.int64)) if len(shape) else 1
    tail = (2,) if bits == 128 else ()
    out = np.empty((max(n, 1),) + tail, dtype=np.uint64)
    s = np.frombuffer(bytes(seed), dtype=np.uint8).copy()
    (_sample_ring64_bits if bits == 64 else _sample_ring128_bits)(_ptr(s), n, policy, _ptr(out))
    return Ring(bits, out[:n].reshape(shape + tail))


def sample_uniform_seeded_bits(shape, seed: bytes, policy: int = 0) -> Bits:
    """SampleSeededOp::bit_kernel (host/ops.rs:2019-2036)."""
    shape = tuple(shape)
    n = int(np.prod(shape, dtype=np.int64)) if len(shape) else 1
    out = np.zeros((n + 7) // 8, dtype=np.uint8)
    s = np.frombuffer(bytes(seed), dtype=np.uint8).copy()
    _sample_bits(_ptr(s), n, policy, _ptr(out))
    return Bits(shape, out)


# ---------------------------------------------------------------------------
# session
# ---------------------------------------------------------------------------

class Session:
    """SyncSession stand-in (execution/synchronous.rs) with explicit entropy."""

    def __init__(self, session_id: str = "b200-bench", entropy_key: bytes = bytes(range(16)),
                 bit_policy: int = BIT_POLICY_STREAM_LSB0):
        self.sid = session_id_bytes(session_id)
        self.entropy = EntropySource(entropy_key)
        self.bit_policy = bit_policy
        self._setup = None

    # host/prim.rs:113-121
    def gen_key(self) -> bytes:
        return self.entropy.next16()

    # host/prim.rs:73-79
    def sync_key_random(self) -> bytes:
        return self.entropy.next16()

    def derive_seed(self, sync_key: bytes, key: bytes) -> bytes:
        return derive_seed(self.sid, sync_key, key)

    # replicated/setup.rs:39-58 via execution/synchronous.rs:297-307 (lazy, cached)
    def setup(self):
        if self._setup is None:
            k00 = self.gen_key()
            k11 = self.gen_key()
            k22 = self.gen_key()
            k10, k21, k02 = k11, k22, k00
            self._setup = [[k00, k10], [k11, k21], [k22, k02]]
        return self._setup

    def sample_like(self, x, seed: bytes):
        """sample_uniform_seeded typed by the operand (ring or bit tensor)."""
        if isinstance(x, Bits):
            return sample_uniform_seeded_bits(x.shape, seed, self.bit_policy)
        return sample_uniform_seeded(x.bits, x.shape, seed)

    def sample_uniform(self, like):
        """SampleOp (host/ops.rs:1883-1957): unseeded = seed drawn from entropy."""
        return self.sample_like(like, self.entropy.next16())


# ---------------------------------------------------------------------------
# replicated protocols.  RepTensor.shares = [[x00,x10],[x11,x21],[x22,x02]]
# ---------------------------------------------------------------------------

class RepTensor:
    def __init__(self, shares):
        self.shares = shares

    def distinct(self):
        """(x0, x1, x2) -- also asserts the two copies of each share agree."""
        (x00, x10), (x11, x21), (x22, x02) = self.shares
        assert x00 == x02 and x10 == x11 and x21 == x22, "inconsistent replicated sharing"
        return x00, x10, x21

    @property
    def shape(self):
        return self.shares[0][0].shape


class AdtTensor:
    def __init__(self, shares):
        self.shares = shares


class MirTensor:
    def __init__(self, values):
        self.values = values


def gen_seeds(sess: Session):
    """SeedsGen::gen_seeds (replicated/zero_share.rs:129-156)."""
    (k00, k10), (k11, k21), (k22, k02) = sess.setup()
    sync_key0 = sess.sync_key_random()
    sync_key1 = sess.sync_key_random()
    sync_key2 = sess.sync_key_random()
    s00 = sess.derive_seed(sync_key0, k00)
    s10 = sess.derive_seed(sync_key1, k10)
    s11 = sess.derive_seed(sync_key1, k11)
    s21 = sess.derive_seed(sync_key2, k21)
    s22 = sess.derive_seed(sync_key2, k22)
    s02 = sess.derive_seed(sync_key0, k02)
    return [[s00, s10], [s11, s21], [s22, s02]]


def gen_zero_share(sess: Session, like):
    """ZeroShareGen::gen_zero_share (replicated/zero_share.rs:82-108)."""
    (s00, s10), (s11, s21), (s22, s02) = gen_seeds(sess)
    r00 = sess.sample_like(like, s00)
    r10 = sess.sample_like(like, s10)
    alpha0 = sub(r00, r10)
    r11 = sess.sample_like(like, s11)
    r21 = sess.sample_like(like, s21)
    alpha1 = sub(r11, r21)
    r22 = sess.sample_like(like, s22)
    r02 = sess.sample_like(like, s02)
    alpha2 = sub(r22, r02)
    return alpha0, alpha1, alpha2


def rep_share(sess: Session, x, owner: int = 0) -> RepTensor:
    """ShareOp::ring_kernel (replicated/convert.rs:49-160). owner in {0,1,2}, or 3 = outsider."""
    (k00, k10), (k11, k21), (k22, k02) = sess.setup()
    if owner == 0:
        sync_key = sess.sync_key_random()
        seed0 = sess.derive_seed(sync_key, k00)
        x00 = sess.sample_like(x, seed0)
        x10 = sub(x, x00)
        seed2 = sess.derive_seed(sync_key, k02)
        x22 = zeros_like(x)
        x02 = sess.sample_like(x, seed2)
        x11 = x10
        x21 = zeros_like(x)
    elif owner == 1:
        sync_key = sess.sync_key_random()
        seed1 = sess.derive_seed(sync_key, k11)
        x11 = sess.sample_like(x, seed1)
        x21 = sub(x, x11)
        seed0 = sess.derive_seed(sync_key, k10)
        x00 = zeros_like(x)
        x10 = sess.sample_like(x, seed0)
        x22 = x21
        x02 = zeros_like(x)
    elif owner == 2:
        sync_key = sess.sync_key_random()
        seed2 = sess.derive_seed(sync_key, k22)
        x22 = sess.sample_like(x, seed2)
        x02 = sub(x, x22)
        seed1 = sess.derive_seed(sync_key, k21)
        x11 = zeros_like(x)
        x21 = sess.sample_like(x, seed1)
        x00 = x02
        x10 = zeros_like(x)
    else:
        sync_key0 = sess.sync_key_random()
        sync_key1 = sess.sync_key_random()
        seed00 = sess.derive_seed(sync_key0, k00)
        seed02 = sess.derive_seed(sync_key0, k02)
        seed11 = sess.derive_seed(sync_key1, k11)
        seed10 = sess.derive_seed(sync_key1, k10)
        x0 = sess.sample_like(x, seed00)
        x1 = sess.sample_like(x, seed11)
        x2 = sub(sub(x, x0), x1)
        x00 = sess.sample_like(x, seed00)
        x10 = sess.sample_like(x, seed10)
        x11 = sess.sample_like(x, seed11)
        x21 = x2
        x22 = x2
        x02 = sess.sample_like(x, seed02)
    return RepTensor([[x00, x10], [x11, x21], [x22, x02]])


def rep_reveal(x: RepTensor, receiver: int = 0):
    """RevealOp::host_ring_kernel (replicated/convert.rs:280-313)."""
    (x00, x10), (x11, x21), (x22, x02) = x.shares
    if receiver == 1:
        return add(add(x02, x11), x21)
    if receiver == 2:
        return add(add(x02, x10), x22)
    return add(add(x00, x10), x21)


def _sharewise2(f, x: RepTensor, y: RepTensor) -> RepTensor:
    return RepTensor([[f(x.shares[i][j], y.shares[i][j]) for j in range(2)] for i in range(3)])


def rep_add(x: RepTensor, y: RepTensor) -> RepTensor:
    """AddOp::rep_rep_kernel (replicated/arith.rs:7-37); XorOp::rep_kernel for bits (:510-523)."""
    return _sharewise2(add, x, y)


def rep_sub(x: RepTensor, y: RepTensor) -> RepTensor:
    """SubOp::rep_rep_kernel (replicated/arith.rs:197-228)."""
    return _sharewise2(sub, x, y)


def rep_neg(x: RepTensor) -> RepTensor:
    """NegOp::rep_rep_kernel (replicated/arith.rs:560-583); bit flavour negates only x00 and x02 (:525-558)."""
    (x00, x10), (x11, x21), (x22, x02) = x.shares
    if isinstance(x00, Bits):
        return RepTensor([[neg(x00), x10], [x11, x21], [x22, neg(x02)]])
    return RepTensor([[neg(v) for v in row] for row in x.shares])


def rep_shl(amount: int, x: RepTensor) -> RepTensor:
    """ShlOp::rep_kernel (replicated/arith.rs:585-609)."""
    return RepTensor([[shl(amount, v) for v in row] for row in x.shares])


def rep_add_mir(x: RepTensor, y: MirTensor) -> RepTensor:
    """AddOp::rep_mir_kernel (replicated/arith.rs:81-121)."""
    y0, y1, y2 = y.values
    (x00, x10), (x11, x21), (x22, x02) = x.shares
    z00 = add(x00, y0)
    z02 = add(x02, y2)
    s0 = z00.shape
    s1 = add(x11, y1).shape
    s2 = z02.shape
    return RepTensor([[z00, broadcast(x10, s0)], [broadcast(x11, s1), broadcast(x21, s1)],
                      [broadcast(x22, s2), z02]])


def rep_mir_add(x: MirTensor, y: RepTensor) -> RepTensor:
    """AddOp::mir_rep_kernel (replicated/arith.rs:39-79)."""
    x0, x1, x2 = x.values
    (y00, y10), (y11, y21), (y22, y02) = y.shares
    z00 = add(x0, y00)
    z02 = add(x2, y02)
    s0, s1, s2 = z00.shape, add(x1, y11).shape, z02.shape
    return RepTensor([[z00, broadcast(y10, s0)], [broadcast(y11, s1), broadcast(y21, s1)],
                      [broadcast(y22, s2), z02]])


def rep_sub_mir(x: RepTensor, y: MirTensor) -> RepTensor:
    """SubOp::rep_mir_kernel (replicated/arith.rs:275-315)."""
    y0, y1, y2 = y.values
    (x00, x10), (x11, x21), (x22, x02) = x.shares
    z00 = sub(x00, y0)
    z02 = sub(x02, y2)
    s0, s1, s2 = z00.shape, sub(x11, y1).shape, z02.shape
    return RepTensor([[z00, broadcast(x10, s0)], [broadcast(x11, s1), broadcast(x21, s1)],
                      [broadcast(x22, s2), z02]])


def rep_mir_sub(x: MirTensor, y: RepTensor) -> RepTensor:
    """SubOp::mir_rep_kernel (replicated/arith.rs:230-273)."""
    x0, x1, x2 = x.values
    (y00, y10), (y11, y21), (y22, y02) = y.shares
    z00 = sub(x0, y00)
    z10, z11, z21, z22 = neg(y10), neg(y11), neg(y21), neg(y22)
    z02 = sub(x2, y02)
    s0, s1, s2 = z00.shape, sub(x1, y11).shape, z02.shape
    return RepTensor([[z00, broadcast(z10, s0)], [broadcast(z11, s1), broadcast(z21, s1)],
                      [broadcast(z22, s2), z02]])


def rep_mul_mir(x: RepTensor, y: MirTensor) -> RepTensor:
    """MulOp::rep_mir_kernel (replicated/arith.rs:402-434)."""
    return RepTensor([[mul(v, y.values[i]) for v in x.shares[i]] for i in range(3)])


def rep_mir_mul(x: MirTensor, y: RepTensor) -> RepTensor:
    """MulOp::mir_rep_kernel (replicated/arith.rs:369-400)."""
    return RepTensor([[mul(x.values[i], v) for v in y.shares[i]] for i in range(3)])


def _rep_bilinear(sess: Session, op, x: RepTensor, y: RepTensor) -> RepTensor:
    (x00, x10), (x11, x21), (x22, x02) = x.shares
    (y00, y10), (y11, y21), (y22, y02) = y.shares
    v0 = add(add(op(x00, y00), op(x00, y10)), op(x10, y00))
    v1 = add(add(op(x11, y11), op(x11, y21)), op(x21, y11))
    v2 = add(add(op(x22, y22), op(x22, y02)), op(x02, y22))
    a0, a1, a2 = gen_zero_share(sess, v0)
    z0, z1, z2 = add(v0, a0), add(v1, a1), add(v2, a2)
    return RepTensor([[z0, z1], [z1, z2], [z2, z0]])


def rep_mul(sess: Session, x: RepTensor, y: RepTensor) -> RepTensor:
    """MulOp::rep_rep_kernel (replicated/arith.rs:317-367); AndOp::rep_kernel for bits (:495-508)."""
    return _rep_bilinear(sess, mul, x, y)


def rep_dot(sess: Session, x: RepTensor, y: RepTensor) -> RepTensor:
    """DotOp::rep_rep_kernel (replicated/arith.rs:436-493)."""
    return _rep_bilinear(sess, dot, x, y)


rep_and = rep_mul
rep_xor = rep_add


# -- additive (2-party) helpers: adt = (player0, player1), provider = player2 --

def rep_to_adt(x: RepTensor) -> AdtTensor:
    """RepToAdtOp::rep_to_adt_kernel, adt=(rep0,rep1) branch (additive/convert.rs:28-37)."""
    (x00, x10), (_x11, x21), _ = x.shares
    return AdtTensor([add(x00, x10), x21])


def adt_reveal(x: AdtTensor):
    """RevealOp::host_adt_kernel (additive/ops.rs:72-84)."""
    return add(x.shares[0], x.shares[1])


def gen_trunc_mask(sess: Session, amount: int, like: Ring):
    """TruncMaskGen::gen_trunc_mask at the provider (additive/trunc.rs:22-65)."""
    r = sess.sample_uniform(like)
    r_msb = shr(like.bits - 1, r)
    r_top = shr(amount + 1, shl(1, r))
    key = sess.gen_key()

    def share(v):
        sync_key = sess.sync_key_random()
        seed = sess.derive_seed(sync_key, key)
        x0 = sess.sample_like(like, seed)
        x1 = sub(v, x0)
        return AdtTensor([x0, x1])

    r_shared = share(r)
    r_top_shared = share(r_top)
    r_msb_shared = share(r_msb)
    return r_shared, r_top_shared, r_msb_shared


def adt_trunc_pr(sess: Session, amount: int, x: AdtTensor) -> AdtTensor:
    """TruncPrProvider::trunc_pr (additive/trunc.rs:114-169)."""
    x0, x1 = x.shares
    bits = x0.bits
    r, r_top, r_msb = gen_trunc_mask(sess, amount, x0)
    k = bits - 1
    ones = fill(bits, 1, x0.shape)
    upshifter = shl(k - 1, ones)
    downshifter = shl(k - amount - 1, ones)
    # adt + host(player0): public value lands on share 0 (additive/ops.rs:105-126)
    x_positive = AdtTensor([add(x0, upshifter), x1])
    masked = AdtTensor([add(x_positive.shares[0], r.shares[0]), add(x_positive.shares[1], r.shares[1])])
    c = adt_reveal(masked)
    c_no_msb = shl(1, c)
    c_top = shr(amount + 1, c_no_msb)
    c_msb = shr(bits - 1, c)
    # overflow = r_msb + c_msb - r_msb*c_msb - r_msb*c_msb
    t = AdtTensor([add(r_msb.shares[0], c_msb), r_msb.shares[1]])
    m = AdtTensor([mul(r_msb.shares[0], c_msb), mul(r_msb.shares[1], c_msb)])
    t = AdtTensor([sub(t.shares[0], m.shares[0]), sub(t.shares[1], m.shares[1])])
    overflow = AdtTensor([sub(t.shares[0], m.shares[0]), sub(t.shares[1], m.shares[1])])
    shifted_overflow = AdtTensor([shl(k - amount, overflow.shares[0]), shl(k - amount, overflow.shares[1])])
    # c_top - r_top: host(player0) - adt (additive/ops.rs:205-229)
    y = AdtTensor([sub(c_top, r_top.shares[0]), neg(r_top.shares[1])])
    y_positive = AdtTensor([add(y.shares[0], shifted_overflow.shares[0]),
                            add(y.shares[1], shifted_overflow.shares[1])])
    return AdtTensor([sub(y_positive.shares[0], downshifter), y_positive.shares[1]])


def adt_to_rep(sess: Session, x: AdtTensor) -> RepTensor:
    """AdtToRepOp::kernel, provider = rep player2 branch (replicated/convert.rs:392-497)."""
    x0, x1 = x.shares
    sync_key0 = sess.sync_key_random()
    sync_key1 = sess.sync_key_random()
    k = sess.gen_key()
    seed1 = sess.derive_seed(sync_key0, k)
    seed2 = sess.derive_seed(sync_key1, k)
    y0 = sess.sample_like(x0, seed1)
    y1 = sess.sample_like(x1, seed2)
    y0_provider = sess.sample_like(x0, seed1)
    y1_provider = sess.sample_like(x0, seed2)
    c = add(sub(x0, y0), sub(x1, y1))
    return RepTensor([[y0, c], [c, y1], [y1_provider, y0_provider]])


def rep_trunc_pr(sess: Session, amount: int, x: RepTensor) -> RepTensor:
    """TruncPrOp::rep_kernel (replicated/fixedpoint.rs:80-103)."""
    x_adt = rep_to_adt(x)
    y_adt = adt_trunc_pr(sess, amount, x_adt)
    return adt_to_rep(sess, y_adt)


# -- bits ---------------------------------------------------------------------

def rep_split(sess: Session, x: RepTensor):
    """PlacementSplit::split (replicated/bits.rs:84-134)."""
    (x00, x10), (_x11, x21), (x22, _x02) = x.shares
    left = add(x00, x10)
    bsl = bit_decompose(left)
    x2_on_1 = bit_decompose(x21)
    x2_on_2 = bit_decompose(x22)
    p0_zero = fill_bits(0, bsl.shape)
    p1_zero = fill_bits(0, x2_on_1.shape)
    p2_zero = fill_bits(0, x2_on_2.shape)
    rep_bsl = rep_share(sess, bsl, owner=0)
    rep_bsr = RepTensor([[p0_zero, p0_zero], [p1_zero, x2_on_1], [x2_on_2, p2_zero]])
    return rep_bsl, rep_bsr


def rep_shl_dim(amount: int, bit_length: int, x: RepTensor) -> RepTensor:
    """ShlDimOp::rep_kernel: share-wise shl_dim (replicated/ops.rs)."""
    return RepTensor([[shl_dim(amount, bit_length, v) for v in row] for row in x.shares])


def binary_adder(sess: Session, x: RepTensor, y: RepTensor, ring_size: int) -> RepTensor:
    """BinaryAdder::binary_adder, Kogge-Stone (replicated/misc.rs:176-242)."""
    log_r = int(math.log2(ring_size))
    g = rep_and(sess, x, y)
    p_store = rep_xor(x, y)
    p = p_store
    for i in range(log_r):
        p1 = rep_shl_dim(1 << i, ring_size, p)
        g1 = rep_shl_dim(1 << i, ring_size, g)
        p_and_g = rep_and(sess, p, g1)
        g = rep_xor(g, p_and_g)
        p = rep_and(sess, p, p1)
    c = rep_shl_dim(1, ring_size, g)
    return rep_xor(c, p_store)


def rep_bit_decompose(sess: Session, x: RepTensor) -> RepTensor:
    """BitDecomposeOp::rep_ring_kernel (replicated/bits.rs:6-20)."""
    bits = x.shares[0][0].bits
    x0, x1 = rep_split(sess, x)
    return binary_adder(sess, x0, x1, bits)


def rep_index(index: int, x: RepTensor) -> RepTensor:
    """IndexOp on RepBitArray: share-wise index_axis(0, index) (replicated/ops.rs)."""
    return RepTensor([[index_axis0(index, v) for v in row] for row in x.shares])


def rep_msb(sess: Session, x: RepTensor) -> RepTensor:
    """MsbOp::rep_bit_kernel (replicated/arith.rs:611-626): full decomposition, plane k-1."""
    bits = x.shares[0][0].bits
    return rep_index(bits - 1, rep_bit_decompose(sess, x))


def gen_dabit(sess: Session, bits: int, shape):
    """DaBitProvider::gen_dabit (additive/dabit.rs:38-69)."""
    like_b = Bits(shape, np.zeros((int(np.prod(shape, dtype=np.int64)) + 7) // 8, np.uint8))
    like_r = fill(bits, 0, shape)
    b = sess.sample_uniform(like_b)
    br = ring_inject(bits, 0, b)
    key = sess.gen_key()
    seed_b = sess.derive_seed(sess.sync_key_random(), key)
    seed_br = sess.derive_seed(sess.sync_key_random(), key)
    b0_provider = sess.sample_like(like_b, seed_b)
    b0 = sess.sample_like(like_b, seed_b)
    b1 = sub(b, b0_provider)
    br0_provider = sess.sample_like(like_r, seed_br)
    br0 = sess.sample_like(like_r, seed_br)
    br1 = sub(br, br0_provider)
    return AdtTensor([br0, br1]), AdtTensor([b0, b1])


def rep_ring_inject(sess: Session, bits: int, bit_idx: int, x: RepTensor) -> RepTensor:
    """RingInjectOp::rep_kernel (replicated/convert.rs:316-390)."""
    (x00, _), _, (x22, _) = x.shares
    b_ring, b_bin = gen_dabit(sess, bits, x22.shape)
    x_adt = rep_to_adt(x)
    c = AdtTensor([add(x_adt.shares[0], b_bin.shares[0]), add(x_adt.shares[1], b_bin.shares[1])])
    c_open = adt_reveal(c)
    c_ring = ring_inject(bits, 0, c_open)
    # b_ring + c_ring - b_ring*c_ring - b_ring*c_ring   (adt op host(player0))
    t = AdtTensor([add(b_ring.shares[0], c_ring), b_ring.shares[1]])
    m = AdtTensor([mul(b_ring.shares[0], c_ring), mul(b_ring.shares[1], c_ring)])
    t = AdtTensor([sub(t.shares[0], m.shares[0]), sub(t.shares[1], m.shares[1])])
    x_adt_ring = AdtTensor([sub(t.shares[0], m.shares[0]), sub(t.shares[1], m.shares[1])])
    shifted = AdtTensor([shl(bit_idx, x_adt_ring.shares[0]), shl(bit_idx, x_adt_ring.shares[1])])
    return adt_to_rep(sess, shifted)


def rep_msb_ring(sess: Session, x: RepTensor) -> RepTensor:
    """MsbOp::rep_ring_kernel (replicated/arith.rs:628-641)."""
    bits = x.shares[0][0].bits
    return rep_ring_inject(sess, bits, 0, rep_msb(sess, x))


def rep_less(sess: Session, x: RepTensor, y: RepTensor) -> RepTensor:
    """LessOp::rep_kernel (replicated/compare.rs:91-104): msb(x - y)."""
    return rep_msb(sess, rep_sub(x, y))


def rep_greater(sess: Session, x: RepTensor, y: RepTensor) -> RepTensor:
    """GreaterOp::rep_kernel (replicated/compare.rs:136-149): msb(y - x)."""
    return rep_msb(sess, rep_sub(y, x))


def rep_mux(sess: Session, s: RepTensor, x: RepTensor, y: RepTensor) -> RepTensor:
    """MuxOp::rep_kernel (replicated/control_flow.rs:9-25): s*(x-y)+y."""
    return rep_add(rep_mul(sess, s, rep_sub(x, y)), y)


def rep_mux_bit_selector(sess: Session, bits: int, s: RepTensor, x: RepTensor, y: RepTensor) -> RepTensor:
    """MuxOp::rep_bit_selector_kernel (replicated/control_flow.rs:27-45): ring_inject(s) then s*(x-y)+y."""
    return rep_mux(sess, rep_ring_inject(sess, bits, 0, s), x, y)


def rep_bit_compose(sess: Session, bits: int, x: RepTensor) -> RepTensor:
    """BitComposeOp::rep_kernel (replicated/bits.rs:23-46): sum_i ring_inject(i, plane i), folded from zeros."""
    planes = [rep_index(i, x) for i in range(bits)]
    acc = rep_fill(bits, 0, planes[0].shares[0][0].shape)
    for i, plane in enumerate(planes):
        acc = rep_add(acc, rep_ring_inject(sess, bits, i, plane))
    return acc


def mir_fill(bits: int, value: int, shape) -> MirTensor:
    v = fill(bits, value, shape)
    return MirTensor([v, v, v])


def rep_maximum(sess: Session, xs: List[RepTensor]) -> RepTensor:
    """MaximumOp::kernel (replicated/softmax.rs:10-53)."""
    n = len(xs)
    if n == 0:
        raise ValueError("maximum op needs a non-empty array of tensors")
    if n == 1:
        return xs[0]
    max1 = rep_maximum(sess, xs[: n // 2])
    max2 = rep_maximum(sess, xs[n // 2:])
    bits = max1.shares[0][0].bits
    lesser = rep_less(sess, max1, max2)
    lesser_ring = rep_ring_inject(sess, bits, 0, lesser)
    ones = mir_fill(bits, 1, lesser_ring.shape)
    # lesser_ring * max2 + (ones - lesser_ring) * max1   (left-to-right evaluation)
    a = rep_mul(sess, lesser_ring, max2)
    b = rep_mul(sess, rep_mir_sub(ones, lesser_ring), max1)
    return rep_add(a, b)


def rep_relu(sess: Session, x: RepTensor) -> RepTensor:
    """ReluOp::rep_ring_kernel (fixedpoint/ops.rs:314-330): mux(msb_ring(x), 0, x)."""
    bits = x.shares[0][0].bits
    sign = rep_msb_ring(sess, x)
    zeros = RepTensor([[zeros_like(v) for v in row] for row in x.shares])
    return rep_mux(sess, sign, zeros, x)


# ---------------------------------------------------------------------------
# exp / pow2 / division / sigmoid (replicated/exp.rs, division.rs, arith.rs:676-743,
# fixedpoint/mod.rs:74-151, misc.rs:5-40).  Fixed-point values are plain RepTensors plus
# (integral, fractional) precision passed explicitly, like RepFixedTensor's fields.
# ---------------------------------------------------------------------------

def fixed_constant(bits: int, value: float, precision: int) -> int:
    """Constant::Fixed -> ring value (kernels/constants.rs:68-72,103-105): ((v * 2^p as f64) as iK) as uK."""
    import struct
    scaled = float(value) * float(1 << precision)
    if scaled != scaled:
        v = 0
    else:
        lim = float(1 << (bits - 1))
        v = (1 << (bits - 1)) - 1 if scaled >= lim else (-(1 << (bits - 1)) if scaled <= -lim else int(scaled))
    return v % (1 << bits)


def rep_fill(bits: int, value: int, shape) -> RepTensor:
    """FillOp::rep_ring64/128_kernel (replicated/ops.rs:91-160): shares [[v,0],[0,0],[0,v]]."""
    v, z = fill(bits, value, shape), fill(bits, 0, shape)
    return RepTensor([[v, z], [z, z], [z, v]])


def rep_fill_bits(value: int, shape) -> RepTensor:
    v, z = fill_bits(value, shape), fill_bits(0, shape)
    return RepTensor([[v, z], [z, z], [z, v]])


def rep_broadcast(x: RepTensor, shape) -> RepTensor:
    return RepTensor([[v if v.shape == tuple(shape) else broadcast(v, shape) for v in row] for row in x.shares])


def prefix_op(sess, xs, op):
    """ReplicatedPlacement::prefix_op (replicated/misc.rs:5-28)."""
    v_len = len(xs)
    log_r = int(math.ceil(math.log2(v_len))) if v_len > 1 else 0
    res = list(xs)
    for i in range(log_r):
        for j in range((2 ** log_r) // (2 ** (i + 1))):
            y = 2 ** i + j * 2 ** (i + 1) - 1
            for k in range(1, 2 ** i + 1):
                if y + k < v_len:
                    res[y + k] = op(sess, res[y], res[y + k])
    return res


def prefix_or(sess, xs):
    """misc.rs:30-40: or(x, y) = xor(xor(x, y), and(x, y))."""
    return prefix_op(sess, xs, lambda s, x, y: rep_xor(rep_xor(x, y), rep_and(s, x, y)))


def prefix_and(sess, xs):
    """misc.rs:42-52 (used only by the reference's tests)."""
    return prefix_op(sess, xs, lambda s, x, y: rep_and(s, x, y))


def sign_from_msb(bits: int, msb_ring: RepTensor) -> RepTensor:
    """SignFromMsb (division.rs:88-104): 1 - 2 msb."""
    double = rep_shl(1, msb_ring)
    return rep_mir_sub(mir_fill(bits, 1, msb_ring.shape), double)


def top_most(sess, max_bits: int, x: RepTensor):
    """TopMost::top_most (division.rs:146-177)."""
    x_bits = rep_bit_decompose(sess, x)
    x_rev = [rep_index(max_bits - i - 1, x_bits) for i in range(max_bits)]
    y = prefix_or(sess, x_rev)
    y.reverse()
    z = [rep_sub(y[i], y[i + 1]) for i in range(max_bits - 1)]
    z.append(y[max_bits - 1])
    return z


def top_most_index(sess, bits: int, max_bits: int, x: RepTensor) -> RepTensor:
    """TopMostIndex::top_most_index (division.rs:183-206)."""
    z = top_most(sess, max_bits, x)
    z_ring = [rep_ring_inject(sess, bits, 0, e) for e in z]
    s_vec = [rep_shl(max_bits - i - 1, zi) for i, zi in enumerate(z_ring)]
    res = rep_shl(0, s_vec[max_bits - 1])
    for item in s_vec[1:max_bits]:
        res = rep_add(res, item)
    return res


def div_norm(sess, bits: int, max_bits: int, x: RepTensor):
    """DivNorm::norm (division.rs:110-139)."""
    msb = rep_msb_ring(sess, x)
    sign = sign_from_msb(bits, msb)
    abs_x = rep_mul(sess, sign, x)
    tm = top_most_index(sess, bits, max_bits, abs_x)
    upshifted = rep_mul(sess, x, tm)
    signed_topmost = rep_mul(sess, sign, tm)
    return upshifted, signed_topmost


def approximate_reciprocal(sess, bits: int, int_precision: int, frac_precision: int, x: RepTensor) -> RepTensor:
    """ApproximateReciprocal (division.rs:208-249)."""
    total = int_precision + frac_precision
    upshifted, signed_topmost = div_norm(sess, bits, total, x)
    alpha = mir_fill(bits, fixed_constant(bits, 2.9142, total), x.shape)
    d = rep_mir_sub(alpha, rep_shl(1, upshifted))
    w = rep_mul(sess, d, signed_topmost)
    return rep_trunc_pr(sess, 2 * int_precision, w)


def rep_div(sess, i_prec: int, f_prec: int, x: RepTensor, y: RepTensor) -> RepTensor:
    """DivOp::rep_rep_kernel (division.rs:5-86), Goldschmidt."""
    bits = x.shares[0][0].bits
    k = i_prec + f_prec
    assert 2 * k <= bits, f"2 * (fractional_precision + integral_precision) = {2 * k}, BitLength = {bits}"
    theta = int(math.ceil(math.log2(k / math.log2(17.0))))
    w = approximate_reciprocal(sess, bits, i_prec, f_prec, y)
    rep_alpha = mir_fill(bits, fixed_constant(bits, 1.0, 2 * f_prec), x.shape)
    init_prod = rep_mul(sess, y, w)
    init_prod_resized = rep_broadcast(init_prod, x.shape)
    a = rep_mir_sub(rep_alpha, init_prod_resized)
    b = rep_mul(sess, x, w)
    b = rep_trunc_pr(sess, f_prec, b)
    for _ in range(theta):
        next_a = rep_mul(sess, a, a)
        next_b = rep_mul(sess, b, rep_mir_add(rep_alpha, a))
        a = rep_trunc_pr(sess, 2 * f_prec, next_a)
        b = rep_trunc_pr(sess, 2 * f_prec, next_b)
    b = rep_mul(sess, b, rep_mir_add(rep_alpha, a))
    return rep_trunc_pr(sess, 2 * f_prec, b)


def p_1045():
    """P_1045 (exp.rs:159-170): ln(2)^i / i! for i < 100, from the exact rational of the f64 power."""
    from fractions import Fraction
    out = []
    for i in range(100):
        ln2i = Fraction(math.pow(math.log(2.0), float(i)))
        out.append(float(ln2i / math.factorial(i)))
    return out


def polynomial_eval(sess, f_prec: int, coeffs, x: RepTensor) -> RepTensor:
    """PolynomialEval::polynomial_eval (fixedpoint/mod.rs:95-151); x has fractional precision f_prec."""
    bits = x.shares[0][0].bits
    degree = len(coeffs) - 1
    for c in reversed(coeffs):
        if c < 2.0 ** (-(f_prec + 1)):
            degree -= 1
        else:
            break
    coeffs_mir = [mir_fill(bits, fixed_constant(bits, c, f_prec), x.shape) for c in coeffs[: degree + 1]]
    x_n = [x for _ in range(degree)]
    x_pre_mul = prefix_op(sess, x_n, lambda s, a, b: rep_trunc_pr(s, f_prec, rep_mul(s, a, b)))
    x_mul_coeffs = [rep_mir_mul(coeffs_mir[i + 1], x_pre_mul[i]) for i in range(len(x_pre_mul))]
    acc = x_mul_coeffs[0]
    for t in x_mul_coeffs[1:]:
        acc = rep_add(acc, t)      # AddNOp::rep_kernel: share-wise add_n
    acc = rep_trunc_pr(sess, f_prec, acc)
    return rep_add_mir(acc, coeffs_mir[0])


def pow2_from_bits(sess, bits: int, xs) -> RepTensor:
    """Pow2FromBits (exp.rs:124-157)."""
    ones = rep_fill(bits, 1, xs[0].shape)
    selectors = []
    for i, bit in enumerate(xs):
        pos = rep_shl(1 << i, bit)
        neg_ = rep_sub(ones, bit)
        selectors.append(rep_add(pos, neg_))
    acc = ones
    for y in selectors:
        acc = rep_mul(sess, acc, y)
    return acc


def exp_from_parts(sess, e_int: RepTensor, e_frac: RepTensor, f: int, k: int) -> RepTensor:
    """ExpFromParts (exp.rs:176-217)."""
    amount = k - 2 - f
    x = rep_shl(amount, e_frac)
    e_approx = polynomial_eval(sess, k - 2, p_1045(), x)
    e_prod = rep_mul(sess, e_int, e_approx)
    return rep_trunc_pr(sess, amount, e_prod)


def rep_pow2(sess, i_prec: int, f_prec: int, x: RepTensor) -> RepTensor:
    """Pow2Op::rep_rep_kernel (exp.rs:11-111)."""
    bits = x.shares[0][0].bits
    bit_len_prec = f_prec + i_prec
    x_bits = rep_bit_decompose(sess, x)
    msb_bit = rep_index(bits - 1, x_bits)
    msb = rep_ring_inject(sess, bits, 0, msb_bit)
    abs_x = rep_mux(sess, msb, rep_neg(x), x)
    absolute_bits = rep_bit_decompose(sess, abs_x)
    x_bits_vec = [rep_index(i, absolute_bits) for i in range(bits)]
    higher_bits = [rep_ring_inject(sess, bits, 0, x_bits_vec[i]) for i in range(f_prec, bits)]
    acc = rep_fill(bits, 0, x.shape)
    for i, item in enumerate(higher_bits):
        acc = rep_add(acc, rep_shl(f_prec + i, item))
    fractional_part = rep_sub(abs_x, acc)
    d = pow2_from_bits(sess, bits, higher_bits[0:i_prec])
    g = exp_from_parts(sess, d, fractional_part, f_prec, bit_len_prec)
    one = rep_fill(bits, 1, x.shape)
    one_fixed = rep_shl(f_prec, one)
    inverse = rep_div(sess, i_prec, f_prec, one_fixed, g)
    return rep_mux(sess, msb, inverse, g)


def rep_exp(sess, i_prec: int, f_prec: int, x: RepTensor) -> RepTensor:
    """ExpOp::rep_rep_kernel (exp.rs:219-244)."""
    bits = x.shares[0][0].bits
    log2e = mir_fill(bits, fixed_constant(bits, math.log2(math.exp(1.0)), f_prec), x.shape)
    shifted_exponent = rep_mir_mul(log2e, x)
    exponent = rep_trunc_pr(sess, f_prec, shifted_exponent)
    return rep_pow2(sess, i_prec, f_prec, exponent)


def rep_sigmoid(sess, i_prec: int, f_prec: int, x: RepTensor) -> RepTensor:
    """SigmoidOp::rep_rep_kernel (arith.rs:676-743)."""
    bits = x.shares[0][0].bits
    ones_rep = rep_fill(bits, fixed_constant(bits, 1.0, f_prec), x.shape)
    zeros_rep = rep_fill(bits, 0, x.shape)
    denominator = rep_add(ones_rep, rep_exp(sess, i_prec, f_prec, rep_neg(x)))
    output = rep_div(sess, i_prec, f_prec, ones_rep, denominator)
    max_val = math.log(2.0 ** (i_prec - 1))
    max_val_rep = rep_fill(bits, fixed_constant(bits, max_val, f_prec), x.shape)
    upper = rep_greater(sess, x, max_val_rep)
    upper_ring = rep_ring_inject(sess, bits, 0, upper)
    upper_wall = rep_mux(sess, upper_ring, ones_rep, output)
    lower = rep_greater(sess, rep_neg(max_val_rep), x)
    lower_ring = rep_ring_inject(sess, bits, 0, lower)
    return rep_mux(sess, lower_ring, zeros_rep, upper_wall)


# -- plaintext-side helpers ---------------------------------------------------

def _signed_ints(x: Ring):
    m = 1 << x.bits
    return [v - m if v >> (x.bits - 1) else v for v in (int(t) for t in x.to_ints().reshape(-1))]


def host_less(x: Ring, y: Ring) -> Bits:
    """LessOp::host_ring64/128_kernel (host/ops.rs:2051-2087): (x - y) as signed < 0."""
    d = sub(x, y)
    return Bits.from_array(np.array([int(v < 0) for v in _signed_ints(d)], dtype=np.uint8).reshape(d.shape))


def host_greater(x: Ring, y: Ring) -> Bits:
    """GreaterOp::host_ring64/128_kernel (host/ops.rs:2125-2161): (x - y) as signed > 0."""
    d = sub(x, y)
    return Bits.from_array(np.array([int(v > 0) for v in _signed_ints(d)], dtype=np.uint8).reshape(d.shape))


def host_mux(s: Bits, x: Ring, y: Ring) -> Ring:
    """MuxOp::host_ring_kernel (host/ops.rs:2227-2252): s * (x - y) + y."""
    return add(mul(ring_inject(x.bits, 0, s), sub(x, y)), y)


def host_sign(x: Ring) -> Ring:
    """SignOp::ring64/128_kernel (host/ops.rs:2482-2506)."""
    return Ring.from_ints(x.bits, [-1 if v < 0 else 1 for v in _signed_ints(x)], x.shape)


def cast_ring128_to_ring64(x: Ring) -> Ring:
    """CastOp::ring_reduction_kernel (host/ops.rs:2277-2288)."""
    return Ring(64, x.data[..., 0].copy())


def host_maximum(xs) -> Ring:
    """MaximumOp::host_ring_kernel (host/ops.rs:1432-1452): unsigned order of Wrapping<T>."""
    vals = [t.to_ints().reshape(-1) for t in xs]
    return Ring.from_ints(xs[0].bits, [max(int(v[i]) for v in vals) for i in range(xs[0].size)], xs[0].shape)


# -- shape / indexing host kernels ---------------------------------------------

def _ring_view(x: Ring) -> np.ndarray:
    """Structured view with one item per ring element, so numpy indexing acts on ring elements."""
    if x.bits == 64:
        return x.data
    return x.data.view(np.dtype([("lo", np.uint64), ("hi", np.uint64)])).reshape(x.shape)


def _ring_from_view(bits: int, v: np.ndarray) -> Ring:
    shape = np.shape(v)
    v = np.ascontiguousarray(v).reshape(shape)  # ascontiguousarray promotes 0-d to 1-d
    if bits == 64:
        return Ring(64, v)
    return Ring(128, v.reshape(-1).view(np.uint64).reshape(shape + (2,)))


def expand_dims_shape(shape, axes):
    """RawShape::extend_singletons (host/mod.rs:83-92) after the descending sort of ExpandDimsOp
    (host/ops.rs:1126-1131): pop() yields the axes in ascending order, each inserted into the shape so far."""
    shape = list(shape)
    for a in sorted(axes):
        shape.insert(a, 1)
    return tuple(shape)


def squeeze_shape(shape, axis: Optional[int]):
    """RawShape::squeeze (host/mod.rs:99-113)."""
    shape = list(shape)
    if axis is None:
        return tuple(s for s in shape if s != 1)
    removed = shape.pop(axis)
    if removed != 1:
        raise ValueError("Cannot select an axis to squeeze out which has size not equal to one")
    return tuple(shape)


def expand_dims(axes, x):
    shape = expand_dims_shape(x.shape, axes)
    return x.reshape(shape) if isinstance(x, Ring) else Bits(shape, x.packed)


def squeeze(axis: Optional[int], x):
    shape = squeeze_shape(x.shape, axis)
    return x.reshape(shape) if isinstance(x, Ring) else Bits(shape, x.packed)


def transpose(x):
    """TransposeOp (host/ops.rs:1270-1299): reversed axes; bits up to rank 3 (bitarray.rs:135-188)."""
    if isinstance(x, Bits):
        if len(x.shape) > 3:
            raise NotImplementedError("tranposing not implemented for 4D tensors or larger yet")
        return Bits.from_array(np.ascontiguousarray(x.to_array().T))
    return _ring_from_view(x.bits, _ring_view(x).T)


def concat(axis: int, xs):
    """ConcatOp::ring_kernel / bit_kernel (host/ops.rs:1228-1266).  The bit kernel appends the flat streams
    whatever the axis and sums shape[axis]."""
    if isinstance(xs[0], Bits):
        flat = np.concatenate([t.to_array().reshape(-1) for t in xs])
        shape = list(xs[0].shape)
        shape[axis] = sum(t.shape[axis] for t in xs)
        return Bits.from_array(flat.reshape(shape))
    return _ring_from_view(xs[0].bits, np.concatenate([_ring_view(t) for t in xs], axis=axis))


def index_axis(axis: int, index: int, x):
    """IndexAxisOp (host/ops.rs:740-783).  Bits: BitArrayRepr::index_axis (bitarray.rs:85-108) copies
    new_dim.size() consecutive bits from offset index * stride[axis]."""
    if isinstance(x, Bits):
        shape = list(x.shape)
        stride = int(np.prod(shape[axis + 1:], dtype=np.int64))
        new_shape = shape[:axis] + shape[axis + 1:]
        size = int(np.prod(new_shape, dtype=np.int64)) if new_shape else 1
        flat = x.to_array().reshape(-1)
        off = index * stride
        return Bits.from_array(flat[off:off + size].reshape(new_shape))
    return _ring_from_view(x.bits, np.take(_ring_view(x), index, axis=axis))


def ndarray_slice(length: int, start: int, end: Optional[int], step: Optional[int]):
    """ndarray::Slice semantics (used by SliceInfo, host/mod.rs:153-165): negative start / end count from
    the back, None end = len, negative step walks the selected range backwards from its last element."""
    step = 1 if step is None else step
    s = start + length if start < 0 else start
    e = length if end is None else (end + length if end < 0 else end)
    if not (0 <= s <= length and 0 <= e <= length) or step == 0:
        raise IndexError("slice out of bounds")
    e = max(e, s)
    idx = list(range(s, e))
    return idx[::step] if step > 0 else idx[::-1][::-step]


def slice_ring(info, x: Ring) -> Ring:
    """SliceOp::host_ring_kernel (host/ops.rs:523-535, 686-704); info = [(start, end, step), ...]."""
    if len(info) != len(x.shape):
        raise ValueError("The input dimension of `info` must match the array to be sliced.")
    v = _ring_view(x)
    for ax, (st, en, sp) in enumerate(info):
        v = np.take(v, ndarray_slice(x.shape[ax], st, en, sp), axis=ax)
    return _ring_from_view(x.bits, v)


def diag(x):
    """DiagOp (host/ops.rs:549-582); bits: lengths 1 and 2 only (bitarray.rs:110-128)."""
    if isinstance(x, Bits):
        n = min(x.shape) if x.shape else 1
        flat = x.to_array().reshape(-1)
        if n == 1:
            return Bits.from_array(flat[:1])
        if n == 2:
            strides = [int(np.prod(x.shape[k + 1:], dtype=np.int64)) for k in range(len(x.shape))]
            return Bits.from_array(np.array([flat[0], flat[sum(strides[:2])]], dtype=np.uint8))
        raise NotImplementedError("todo!() in BitArrayRepr::into_diag")
    v = _ring_view(x)
    n = min(x.shape) if x.shape else 1
    idx = tuple(np.arange(n) for _ in x.shape)
    return _ring_from_view(x.bits, v[idx] if x.shape else v.reshape(1))


def select(axis: int, index: Bits, x: Ring) -> Ring:
    """SelectOp::host_ring_kernel (host/ops.rs:662-684, 803-818)."""
    picks = np.nonzero(index.to_array().reshape(-1))[0]
    return _ring_from_view(x.bits, np.take(_ring_view(x), picks, axis=axis))


def ring_div(x: Ring, y: Ring) -> Ring:
    """DivOp::ring_kernel (host/ops.rs:332-343): unsigned quotient, ndarray broadcasting."""
    xb, yb, shape = _bcast(x, y)
    a, b = xb.to_ints().reshape(-1), yb.to_ints().reshape(-1)
    if any(int(v) == 0 for v in b):
        raise ZeroDivisionError("attempt to divide by zero")
    return Ring.from_ints(x.bits, [int(p) // int(q) for p, q in zip(a, b)], shape)


def bit_decompose_ring(x: Ring) -> Ring:
    """BitDecomposeOp::host_ring64/128_kernel (host/ops.rs:856-894): [k, shape...] ring planes."""
    vals = [int(v) for v in x.to_ints().reshape(-1)]
    planes = [[(v >> i) & 1 for v in vals] for i in range(x.bits)]
    return Ring.from_ints(x.bits, planes, (x.bits,) + tuple(x.shape))


def fixedpoint_mean(axis: Optional[int], scaling_base: int, scaling_exp: int, x: Ring) -> Ring:
    """RingFixedpointMeanOp (host/fixedpoint.rs:41-113): sum(axis) * encode(1 / len, base^exp)."""
    length = x.shape[axis] if axis is not None else x.size
    scale = (scaling_base ** scaling_exp) % (1 << x.bits)
    up = (1.0 / float(length)) * float(scale)
    lim = float(1 << (x.bits - 1))
    w = (1 << (x.bits - 1)) - 1 if up >= lim else (-(1 << (x.bits - 1)) if up <= -lim else int(up))
    s = sum_axis(x, axis)
    return mul(s, Ring.from_ints(x.bits, [w] * s.size, s.shape))


def fixedpoint_argmax(axis: int, x: Ring) -> Ring:
    """RingFixedpointArgmaxOp (host/ops.rs:2375-2437): first strict maximum of the signed values."""
    signed = np.array(_signed_ints(x), dtype=object).reshape(x.shape)
    moved = np.moveaxis(signed, axis, 0)
    best = moved[0].copy() if moved.ndim > 1 else moved[0]
    out_shape = moved.shape[1:]
    flat_best = np.array(best, dtype=object).reshape(-1)
    arg = [0] * flat_best.size
    for j in range(moved.shape[0]):
        row = np.array(moved[j], dtype=object).reshape(-1)
        for t in range(flat_best.size):
            if flat_best[t] < row[t]:
                flat_best[t] = row[t]
                arg[t] = j
    return Ring.from_ints(64, arg, out_shape)


# -- replicated share-wise shape ops, softmax, argmax ------------------------------

def _sharewise(f, x: RepTensor) -> RepTensor:
    return RepTensor([[f(v) for v in row] for row in x.shares])


def rep_index_axis(axis: int, index: int, x: RepTensor) -> RepTensor:
    """IndexAxisOp::rep_kernel (replicated/ops.rs:331-360)."""
    return _sharewise(lambda v: index_axis(axis, index, v), x)


def rep_expand_dims(axes, x: RepTensor) -> RepTensor:
    """ExpandDimsOp::rep_kernel (replicated/ops.rs:267-295)."""
    return _sharewise(lambda v: expand_dims(axes, v), x)


def rep_squeeze(axis, x: RepTensor) -> RepTensor:
    """SqueezeOp::rep_kernel (replicated/ops.rs:578-620)."""
    return _sharewise(lambda v: squeeze(axis, v), x)


def rep_transpose(x: RepTensor) -> RepTensor:
    """TransposeOp::rep_kernel (replicated/ops.rs:622-650)."""
    return _sharewise(transpose, x)


def rep_sum(axis, x: RepTensor) -> RepTensor:
    """SumOp::rep_kernel (replicated/arith.rs:167-195)."""
    return _sharewise(lambda v: sum_axis(v, axis), x)


def rep_concat(axis: int, xs) -> RepTensor:
    """ConcatOp::rep_kernel (replicated/ops.rs:48-89)."""
    return RepTensor([[concat(axis, [t.shares[i][j] for t in xs]) for j in range(2)] for i in range(3)])


def rep_cast_64(x: RepTensor) -> RepTensor:
    """CastOp::rep_reduction_kernel (replicated/convert.rs:499-530): share-wise mod 2^64."""
    return _sharewise(cast_ring128_to_ring64, x)


def rep_fixedpoint_mean(axis, scaling_base: int, scaling_exp: int, x: RepTensor) -> RepTensor:
    """RingFixedpointMeanOp::rep_kernel (replicated/fixedpoint.rs:49-78)."""
    return _sharewise(lambda v: fixedpoint_mean(axis, scaling_base, scaling_exp, v), x)


def rep_softmax(sess, axis: int, upmost_index: int, i_prec: int, f_prec: int, x: RepTensor) -> RepTensor:
    """SoftmaxOp::rep_fixed_kernel (replicated/softmax.rs:55-127)."""
    import math
    bits = x.shares[0][0].bits
    xs = [rep_index_axis(axis, i, x) for i in range(upmost_index)]
    xmax = rep_expand_dims([axis], rep_maximum(sess, xs))
    max_diff = rep_sub(x, xmax)
    e_x = rep_exp(sess, i_prec, f_prec, max_diff)
    zeros_rep = rep_fill(bits, 0, e_x.shape)
    min_val = -1.0 * math.log(2.0 ** (i_prec - 1))
    lower_bound = rep_fill(bits, fixed_constant(bits, min_val, f_prec), max_diff.shape)
    threshold = rep_greater(sess, lower_bound, max_diff)
    threshold_ring = rep_ring_inject(sess, bits, 0, threshold)
    normalized = rep_mux(sess, threshold_ring, zeros_rep, e_x)
    e_x_sum = rep_expand_dims([axis], rep_sum(axis, normalized))
    return rep_div(sess, i_prec, f_prec, normalized, e_x_sum)


def rep_argmax(sess, axis: int, upmost_index: int, x: RepTensor) -> RepTensor:
    """RingFixedpointArgmaxOp::rep_ring_kernel + TreeReduceArgmax (replicated/argmax.rs:10-92); result Ring64."""
    bits = x.shares[0][0].bits
    xs = [rep_index_axis(axis, i, x) for i in range(upmost_index)]
    pairs = [(rep_fill(bits, i, item.shape), item) for i, item in enumerate(xs)]

    def combine(a, b):
        comp_ring = rep_ring_inject(sess, bits, 0, rep_less(sess, a[1], b[1]))
        return rep_mux(sess, comp_ring, b[0], a[0]), rep_mux(sess, comp_ring, b[1], a[1])

    def tree(v):
        if len(v) == 1:
            return v[0]
        return combine(tree(v[: len(v) // 2]), tree(v[len(v) // 2:]))

    secret_index, _ = tree(pairs)
    return rep_cast_64(secret_index) if bits == 128 else secret_index


def rep_abs(sess, x: RepTensor) -> RepTensor:
    """AbsOp::rep_ring_kernel (replicated/arith.rs:655-674)."""
    bits = x.shares[0][0].bits
    msb_ring = rep_msb_ring(sess, x)
    return rep_mul(sess, sign_from_msb(bits, msb_ring), x)


def rep_add_n(xs) -> RepTensor:
    """AddNOp::rep_kernel (replicated/arith.rs:123-165)."""
    return RepTensor([[add_n([t.shares[i][j] for t in xs]) for j in range(2)] for i in range(3)])


# -- equality, int2fl, log2 / ln / sqrt (replicated/compare.rs, log.rs, sqrt.rs) -------------------------

P_2524 = [-2.05466671951, -8.8626599391, 6.10585199015, 4.81147460989]   # log.rs:10
Q_2524 = [0.353553425277, 4.54517087629, 6.42784209029, 1.0]             # log.rs:11


def rep_not_bits(x: RepTensor) -> RepTensor:
    """xor(ones_mirrored, x) on shared bits (Z2 add of a public value touches share 0 only, arith.rs:39-79)."""
    (x00, x10), (x11, x21), (x22, x02) = x.shares
    return RepTensor([[neg(x00), x10], [x11, x21], [x22, neg(x02)]])


def rep_equal_zero(sess, x_bits: RepTensor, nbits: int) -> RepTensor:
    """EqualZeroOp::bitdec_bit_kernel (compare.rs:53-72): AND-tree over the negated bits."""
    v_not = [rep_not_bits(rep_index(i, x_bits)) for i in range(nbits)]

    def tree(v):
        if len(v) == 1:
            return v[0]
        a, b = tree(v[: len(v) // 2]), tree(v[len(v) // 2:])
        return rep_and(sess, a, b)

    return tree(v_not)


def rep_equal(sess, x: RepTensor, y: RepTensor) -> RepTensor:
    """EqualOp::rep_kernel (compare.rs:24-40): equal_zero(bit_decompose(x - y))."""
    bits = x.shares[0][0].bits
    return rep_equal_zero(sess, rep_bit_decompose(sess, rep_sub(x, y)), bits)


def int2fl(sess, x: RepTensor, max_bit_len: int, f_prec: int):
    """Int2FL::int2fl (log.rs:166-219): v, p, s, z with (1 - 2s)(1 - z) v 2^p = x."""
    bits = x.shares[0][0].bits
    lam = max_bit_len - 1
    x_bits = rep_bit_decompose(sess, x)
    sign = rep_ring_inject(sess, bits, 0, rep_index(bits - 1, x_bits))            # MsbOp::rep_bit_dec_kernel
    is_zero = rep_ring_inject(sess, bits, 0, rep_equal_zero(sess, x_bits, bits))  # EqualZeroOp::bitdec_ring_kernel
    x_pos = rep_mux(sess, sign, rep_neg(x), x)
    x_pos_bits = rep_bit_decompose(sess, x_pos)
    rev = [rep_index(i, x_pos_bits) for i in range(lam)][::-1]
    b = prefix_or(sess, rev)
    b_ring = [rep_ring_inject(sess, bits, 0, b[i]) for i in range(lam)]
    ones = mir_fill(bits, 1, x.shape)
    zero = rep_fill(bits, 0, x.shape)
    neg_b_sum = zero
    for i in range(lam):
        neg_b_sum = rep_add(neg_b_sum, rep_shl(i, rep_mir_sub(ones, b_ring[i])))
    x_upshifted = rep_mul(sess, x_pos, rep_mir_add(ones, neg_b_sum))
    x_norm = rep_trunc_pr(sess, max_bit_len - 1 - f_prec, x_upshifted)
    bit_count = zero
    for i in range(lam):
        bit_count = rep_add(bit_count, b_ring[i])
    ften = mir_fill(bits, f_prec, x.shape)
    p_res = rep_mul(sess, rep_sub_mir(bit_count, ften), rep_mir_sub(ones, is_zero))
    return x_norm, p_res, sign, is_zero


def rep_log2(sess, i_prec: int, f_prec: int, x: RepTensor) -> RepTensor:
    """Log2Op::rep_rep_kernel (log.rs:13-82)."""
    v, p, _s, _z = int2fl(sess, x, i_prec + f_prec, f_prec)
    p2524 = polynomial_eval(sess, f_prec, P_2524, v)
    q2524 = polynomial_eval(sess, f_prec, Q_2524, v)
    quotient = rep_div(sess, i_prec, f_prec, p2524, q2524)
    return rep_add(rep_shl(f_prec, p), quotient)


def rep_log(sess, i_prec: int, f_prec: int, x: RepTensor) -> RepTensor:
    """LogOp::rep_rep_kernel (log.rs:84-111): ln 2 * log2(x)."""
    bits = x.shares[0][0].bits
    ln2 = mir_fill(bits, fixed_constant(bits, math.log(2.0), f_prec), x.shape)
    return rep_trunc_pr(sess, f_prec, rep_mir_mul(ln2, rep_log2(sess, i_prec, f_prec, x)))


def rep_sqrt(sess, i_prec: int, f_prec: int, x: RepTensor) -> RepTensor:
    """SqrtOp::rep_rep_kernel (sqrt.rs:7-34): 2^(0.5 log2 x)."""
    bits = x.shares[0][0].bits
    log2_x = rep_log2(sess, i_prec, f_prec, x)
    half = mir_fill(bits, fixed_constant(bits, 0.5, f_prec), x.shape)
    exponent = rep_trunc_pr(sess, f_prec, rep_mir_mul(half, log2_x))
    return rep_pow2(sess, i_prec, f_prec, exponent)
