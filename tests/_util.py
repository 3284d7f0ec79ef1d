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
