"""Shared helpers for the tests (not collected)."""
import ctypes

import numpy as np


def has_gpu() -> bool:
    try:
        cuda = ctypes.CDLL("libcuda.so.1")
        if cuda.cuInit(0) != 0:
            return False
        n = ctypes.c_int(0)
        cuda.cuDeviceGetCount(ctypes.byref(n))
        return n.value > 0
    except Exception:
        return False


HAS_GPU = has_gpu()
ENTROPY_KEY = bytes(range(16))
SESSION_ID = "b200-parity"


def rand_ring(rng, bits, shape):
    """Uniform ring elements as uint64[shape] / uint64[shape+(2,)]."""
    shape = tuple(shape)
    tail = (2,) if bits == 128 else ()
    n = int(np.prod(shape + tail, dtype=np.int64))
    return np.frombuffer(rng.bytes(8 * n), dtype=np.uint64).reshape(shape + tail).copy()


def same_ring(gpu_tensor, oracle_ring) -> bool:
    return gpu_tensor.shape == oracle_ring.shape and np.array_equal(gpu_tensor.to_host(), oracle_ring.data)


def same_bits(gpu_tensor, oracle_bits) -> bool:
    return gpu_tensor.shape == oracle_bits.shape and np.array_equal(gpu_tensor.to_host(), oracle_bits.packed)


def same(gpu_tensor, oracle_val) -> bool:
    from oracle import moose_oracle as mo
    if isinstance(oracle_val, mo.Bits):
        return same_bits(gpu_tensor, oracle_val)
    return same_ring(gpu_tensor, oracle_val)


def same_rep(gpu_rep, oracle_rep) -> bool:
    """Share-level parity: the three distinct shares equal the oracle's six (which must be consistent)."""
    o = oracle_rep.distinct()
    return all(same(gpu_rep.share(i), o[i]) for i in range(3))


# A host-dialect computation written for these tests in Moose's textual grammar: the shape / indexing /
# plaintext-helper rows of the kernel table in one graph (values in fixed(.,16) on Ring64).
HOST_TAIL_GRAPH = """
x = Constant{value = HostFloat64Tensor([[1.5, -2.0, 3.25], [0.5, 4.0, -1.0]])}: () -> HostFloat64Tensor () @Host(alice)
xe = RingFixedpointEncode{scaling_base = 2, scaling_exp = 16}: (HostFloat64Tensor) -> HostRing64Tensor (x) @Host(alice)
xt = Transpose: (HostRing64Tensor) -> HostRing64Tensor (xe) @Host(alice)
col0 = IndexAxis{axis = 1, index = 0}: (HostRing64Tensor) -> HostRing64Tensor (xe) @Host(alice)
col = ExpandDims{axis = [1]}: (HostRing64Tensor) -> HostRing64Tensor (col0) @Host(alice)
cat = Concat{axis = 1}: [HostRing64Tensor] -> HostRing64Tensor (col, xe) @Host(alice)
sq = Squeeze{axis = 1}: (HostRing64Tensor) -> HostRing64Tensor (col) @Host(alice)
mean = RingFixedpointMean{axis = 0, scaling_base = 2, scaling_exp = 16}: (HostRing64Tensor) -> HostRing64Tensor (cat) @Host(alice)
dec = RingFixedpointDecode{scaling_base = 2, scaling_exp = 32}: (HostRing64Tensor) -> HostFloat64Tensor (mean) @Host(alice)
amax = RingFixedpointArgmax{axis = 1, upmost_index = 4}: (HostRing64Tensor) -> HostRing64Tensor (cat) @Host(alice)
sl = Slice{slice = {start = 1, end = 4, step = 2}}: (HostRing64Tensor) -> HostRing64Tensor (mean) @Host(alice)
rev = Slice{slice = {start = 0, step = -1}}: (HostRing64Tensor) -> HostRing64Tensor (mean) @Host(alice)
ng = Neg: (HostRing64Tensor) -> HostRing64Tensor (xe) @Host(alice)
lt = Less: (HostRing64Tensor, HostRing64Tensor) -> HostBitTensor (xe, ng) @Host(alice)
gt = Greater: (HostRing64Tensor, HostRing64Tensor) -> HostBitTensor (xe, ng) @Host(alice)
nlt = Neg: (HostBitTensor) -> HostBitTensor (lt) @Host(alice)
either = Or: (HostBitTensor, HostBitTensor) -> HostBitTensor (lt, gt) @Host(alice)
ab = Mux: (HostBitTensor, HostRing64Tensor, HostRing64Tensor) -> HostRing64Tensor (lt, ng, xe) @Host(alice)
sgn = Sign: (HostRing64Tensor) -> HostRing64Tensor (xe) @Host(alice)
mx = Maximum: [HostRing64Tensor] -> HostRing64Tensor (ab, xe) @Host(alice)
dg = Diag: (HostRing64Tensor) -> HostRing64Tensor (xt) @Host(alice)
q = Div: (HostRing64Tensor, HostRing64Tensor) -> HostRing64Tensor (ab, col) @Host(alice)
planes = BitDecompose: (HostRing64Tensor) -> HostRing64Tensor (sq) @Host(alice)
bits = BitDecompose: (HostRing64Tensor) -> HostBitTensor (sq) @Host(alice)
bt = Transpose: (HostBitTensor) -> HostBitTensor (bits) @Host(alice)
b7 = IndexAxis{axis = 0, index = 15}: (HostBitTensor) -> HostBitTensor (bits) @Host(alice)
sel = Select{axis = 1}: (HostRing64Tensor, HostBitTensor) -> HostRing64Tensor (xt, b7) @Host(alice)
shp = Shape: (HostRing64Tensor) -> HostShape (cat) @Host(alice)
tail = Slice{slice = {start = 1, end = 2}}: (HostShape) -> HostShape (shp) @Host(alice)
total = Sum: (HostRing64Tensor) -> HostRing64Tensor (ab) @Host(alice)
absdec = RingFixedpointDecode{scaling_base = 2, scaling_exp = 16}: (HostRing64Tensor) -> HostFloat64Tensor (ab) @Host(alice)
o1 = Output{tag = "mean"}: (HostFloat64Tensor) -> HostFloat64Tensor (dec) @Host(alice)
o2 = Output{tag = "abs"}: (HostFloat64Tensor) -> HostFloat64Tensor (absdec) @Host(alice)
"""
