"""moose_b200 -- host-side mirror of Moose's HostPlacement / ReplicatedPlacement
kernel interface over libmoose_b200.so (hand-written sm_100a CUDA behind a C ABI).

The names and argument meaning follow the reference's placement traits
(moose/src/kernels/*.rs): `HostPlacement.add/sub/mul/dot/shl/shr/neg/
sample_uniform_seeded/derive_seed/gen_key/fill/bit_decompose/ring_inject/...`,
`ReplicatedPlacement.share/reveal/add/sub/mul/dot/trunc_pr/msb/...`.  Every
method is a thin ctypes call into the C ABI declared in include/moose_b200.h;
there is no Python or CPU arithmetic path: importing works anywhere (so the ABI
can be inspected), but creating a `Session` without the built library or without
a CUDA device raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmoose_b200.so")

RING64, RING128, BIT, FLOAT64 = 0, 1, 2, 3
_DT_NAMES = {RING64: "ring64", RING128: "ring128", BIT: "bit", FLOAT64: "float64"}


class MooseB200Error(RuntimeError):
    """Maps to moose::Error (error.rs:8-59); `.code` is the mb_status."""

    def __init__(self, code: int, msg: str):
        super().__init__(f"[{code}] {msg}")
        self.code = code


_lib = None


def lib() -> ctypes.CDLL:
    """Loads libmoose_b200.so; fails loudly when the CUDA extension is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MooseB200Error(-6, f"{LIB_PATH} is missing: run `python -m moose_b200.build` "
                                     "(there is no CPU fallback)")
        _lib = ctypes.CDLL(LIB_PATH)
        _declare(_lib)
    return _lib


_P = ctypes.c_void_p
_PP = ctypes.POINTER(ctypes.c_void_p)
_I = ctypes.c_int
_I64 = ctypes.c_int64
_U64 = ctypes.c_uint64
_U32 = ctypes.c_uint32
_I64P = ctypes.POINTER(ctypes.c_int64)
_U8P = ctypes.c_char_p

# name -> argtypes (restype int unless listed in _RESTYPES)
ABI = {
    "mb_ctx_create": [_I, _PP], "mb_ctx_destroy": [_P], "mb_last_error": [_P],
    "mb_ctx_set_session": [_P, ctypes.c_char_p], "mb_ctx_set_session_raw": [_P, _U8P],
    "mb_ctx_get_session_raw": [_P, _P], "mb_ctx_set_entropy_key": [_P, _U8P], "mb_ctx_sync": [_P],
    "mb_ctx_stream": [_P], "mb_ctx_timer_start": [_P], "mb_ctx_timer_stop": [_P, ctypes.POINTER(ctypes.c_float)],
    "mb_ctx_launch_count": [_P], "mb_ctx_set_gemm_mode": [_P, _I], "mb_ctx_set_fused_bits": [_P, _I],
    "mb_ctx_profile_enable": [_P, _I],
    "mb_ctx_profile_read": [_P, _I, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_uint64)],
    "mb_sample_seeded_at": [_P, _I, _I, _I64P, _U8P, _U64, _PP],
    "mb_ctx_reserve_sms": [_P, _I], "mb_ctx_trace_enable": [_P, _I], "mb_ctx_trace_read": [_P, _P, ctypes.c_size_t, ctypes.POINTER(ctypes.c_size_t)],
    "mb_host_alloc_pinned": [_P, ctypes.c_size_t, _PP], "mb_host_free_pinned": [_P, _P],
    "mb_tensor_from_host": [_P, _I, _I, _I64P, _P, _PP], "mb_tensor_alloc": [_P, _I, _I, _I64P, _PP],
    "mb_tensor_to_host": [_P, _P, _P, ctypes.c_size_t], "mb_tensor_free": [_P, _P],
    "mb_tensor_from_host_async": [_P, _I, _I, _I64P, _P, _PP], "mb_tensor_to_host_async": [_P, _P, _P, ctypes.c_size_t],
    "mb_ctx_wait": [_P, _P],
    "mb_tensor_copy_to_device": [_P, _P, _P], "mb_tensor_copy_from_device": [_P, _P, _P],
    "mb_tensor_dtype": [_P], "mb_tensor_rank": [_P], "mb_tensor_shape": [_P], "mb_tensor_numel": [_P],
    "mb_tensor_nbytes": [_P], "mb_tensor_device_ptr": [_P], "mb_reshape": [_P, _P, _I, _I64P, _PP],
    "mb_prf_key_gen": [_P, _P], "mb_sync_key_random": [_P, _P], "mb_derive_seed": [_P, _U8P, _U8P, _P],
    "mb_sample_seeded": [_P, _I, _I, _I64P, _U8P, _I64, _PP], "mb_sample": [_P, _I, _I, _I64P, _I64, _PP],
    "mb_fill": [_P, _I, _I, _I64P, _U64, _U64, _PP],
    "mb_add": [_P, _P, _P, _PP], "mb_sub": [_P, _P, _P, _PP], "mb_mul": [_P, _P, _P, _PP],
    "mb_neg": [_P, _P, _PP], "mb_shl": [_P, _U64, _P, _PP], "mb_shr": [_P, _U64, _P, _PP],
    "mb_dot": [_P, _P, _P, _PP], "mb_dot_noalloc": [_P, _I, _P, _P, _P, _I64, _I64, _I64], "mb_add_n": [_P, _PP, _I, _PP], "mb_sum": [_P, _P, _I, _PP],
    "mb_broadcast": [_P, _P, _I, _I64P, _PP],
    "mb_fixedpoint_encode": [_P, _I, _U32, _P, _PP], "mb_fixedpoint_decode": [_P, _U32, _P, _PP],
    "mb_less": [_P, _P, _P, _PP], "mb_greater": [_P, _P, _P, _PP], "mb_mux": [_P, _P, _P, _P, _PP],
    "mb_sign": [_P, _P, _PP], "mb_cast_ring128_to_ring64": [_P, _P, _PP], "mb_maximum": [_P, _PP, _I, _PP],
    "mb_xor": [_P, _P, _P, _PP], "mb_and": [_P, _P, _P, _PP], "mb_or": [_P, _P, _P, _PP],
    "mb_bit_decompose": [_P, _P, _PP], "mb_bit_extract": [_P, _U64, _P, _PP],
    "mb_ring_inject": [_P, _I, _U64, _P, _PP], "mb_shl_dim": [_P, _U64, _U64, _P, _PP],
    "mb_index_axis0": [_P, _I64, _P, _PP], "mb_index_axis": [_P, _I, _I64, _P, _PP],
    "mb_expand_dims": [_P, _I64P, _I, _P, _PP], "mb_squeeze": [_P, _I, _P, _PP], "mb_transpose": [_P, _P, _PP],
    "mb_concat": [_P, _I, _PP, _I, _PP], "mb_slice": [_P, _I, _I64P, _I64P, _U8P, _I64P, _P, _PP],
    "mb_diag": [_P, _P, _PP], "mb_select": [_P, _I, _P, _P, _PP], "mb_div": [_P, _P, _P, _PP],
    "mb_bit_decompose_ring": [_P, _P, _PP], "mb_fixedpoint_mean": [_P, _I, _U64, _U32, _P, _PP],
    "mb_fixedpoint_argmax": [_P, _I, _P, _PP],
    "mb_rep_from_shares": [_P, _P, _P, _P, _PP], "mb_rep_share_tensor": [_P, _I], "mb_rep_free": [_P, _P],
    "mb_rep_share_deferred": [_P, _P, _I, _PP], "mb_rep_materialize": [_P, _P],
    "mb_rep_setup": [_P], "mb_rep_share": [_P, _P, _I, _PP], "mb_rep_reveal": [_P, _P, _I, _PP],
    "mb_rep_add": [_P, _P, _P, _PP], "mb_rep_sub": [_P, _P, _P, _PP], "mb_rep_neg": [_P, _P, _PP],
    "mb_rep_shl": [_P, _U64, _P, _PP],
    "mb_rep_add_public": [_P, _P, _P, _PP], "mb_rep_sub_public": [_P, _P, _P, _PP],
    "mb_rep_public_sub": [_P, _P, _P, _PP], "mb_rep_mul_public": [_P, _P, _P, _PP],
    "mb_rep_mul": [_P, _P, _P, _PP], "mb_rep_dot": [_P, _P, _P, _PP], "mb_rep_trunc_pr": [_P, _U32, _P, _PP],
    "mb_rep_bit_decompose": [_P, _P, _PP], "mb_rep_msb": [_P, _P, _PP],
    "mb_rep_ring_inject": [_P, _I, _U64, _P, _PP], "mb_rep_mux": [_P, _P, _P, _P, _PP],
    "mb_rep_bit_compose": [_P, _I, _P, _PP],
    "mb_rep_less": [_P, _P, _P, _PP], "mb_rep_maximum": [_P, _PP, _I, _PP], "mb_rep_relu": [_P, _P, _PP],
    "mb_rep_fill": [_P, _I, _I, _I64P, _U64, _U64, _PP], "mb_rep_greater": [_P, _P, _P, _PP],
    "mb_rep_div": [_P, _U32, _U32, _P, _P, _PP], "mb_rep_pow2": [_P, _U32, _U32, _P, _PP],
    "mb_rep_exp": [_P, _U32, _U32, _P, _PP], "mb_rep_sigmoid": [_P, _U32, _U32, _P, _PP],
    "mb_rep_index_axis": [_P, _I, _I64, _P, _PP], "mb_rep_expand_dims": [_P, _I64P, _I, _P, _PP],
    "mb_rep_squeeze": [_P, _I, _P, _PP], "mb_rep_transpose": [_P, _P, _PP], "mb_rep_reshape": [_P, _P, _I, _I64P, _PP],
    "mb_rep_broadcast": [_P, _P, _I, _I64P, _PP], "mb_rep_slice": [_P, _I, _I64P, _I64P, _U8P, _I64P, _P, _PP],
    "mb_rep_diag": [_P, _P, _PP], "mb_rep_concat": [_P, _I, _PP, _I, _PP], "mb_rep_add_n": [_P, _PP, _I, _PP],
    "mb_rep_sum": [_P, _P, _I, _PP], "mb_rep_cast_ring128_to_ring64": [_P, _P, _PP],
    "mb_rep_fixedpoint_mean": [_P, _I, _U64, _U32, _P, _PP], "mb_rep_abs": [_P, _P, _PP],
    "mb_rep_dot_concat": [_P, _PP, _I, _P, _PP], "mb_rep_fixed_dot": [_P, _U32, _P, _P, _PP],
    "mb_rep_fixed_dot_concat": [_P, _U32, _PP, _I, _P, _PP], "mb_rep_fixed_mul": [_P, _U32, _P, _P, _PP], "mb_rep_equal": [_P, _P, _P, _PP], "mb_rep_log2": [_P, _U32, _U32, _P, _PP], "mb_rep_log": [_P, _U32, _U32, _P, _PP],
    "mb_rep_sqrt": [_P, _U32, _U32, _P, _PP],
    "mb_rep_softmax": [_P, _I, _I, _U32, _U32, _P, _PP], "mb_rep_argmax": [_P, _I, _I, _P, _PP],
    "mb_load_npy": [_P, ctypes.c_char_p, _PP], "mb_load_npy_fixed": [_P, ctypes.c_char_p, _I, _U32, _PP],
    "mb_save_npy": [_P, _P, ctypes.c_char_p], "mb_load_csv": [_P, ctypes.c_char_p, _PP, _I, _PP],
    "mb_save_csv": [_P, _P, ctypes.c_char_p],
    "mb_nccl_unique_id": [_P], "mb_nccl_transport_create": [_P, _P, _I, _I, ctypes.POINTER(ctypes.c_char_p), _PP],
    "mb_nccl_transport_fn": [], "mb_nccl_transport_free": [_P],
    "mb_peer_transport_create": [_P, _I, _I, ctypes.POINTER(ctypes.c_char_p), ctypes.c_size_t, _PP],
    "mb_peer_transport_handle": [_P, _P], "mb_peer_transport_connect": [_P, _P],
    "mb_peer_transport_fn": [], "mb_peer_transport_free": [_P],
    "mb_party_create": [_P, _I, _P, _P, ctypes.POINTER(ctypes.c_char_p), _PP], "mb_party_free": [_P],
    "mb_party_flush": [_P], "mb_party_bytes_sent": [_P], "mb_party_stat": [_P, _I], "mb_pshare_free": [_P, _P], "mb_pshare_tensor": [_P, _I],
    "mb_party_share": [_P, _I, _P, _I, _I, _I64P, _PP], "mb_party_reveal": [_P, _I, _P, _PP],
    "mb_party_add": [_P, _P, _P, _PP], "mb_party_sub": [_P, _P, _P, _PP], "mb_party_add_public": [_P, _P, _P, _PP],
    "mb_party_mul": [_P, _P, _P, _PP], "mb_party_dot": [_P, _P, _P, _PP], "mb_party_trunc_pr": [_P, _U32, _P, _PP],
    "mb_party_msb": [_P, _P, _PP], "mb_party_ring_inject": [_P, _I, _U64, _P, _PP], "mb_party_less": [_P, _P, _P, _PP],
    "mb_party_maximum": [_P, _PP, _I, _PP],
    "mb_graph_parse": [_P, ctypes.c_char_p, _PP], "mb_graph_free": [_P], "mb_graph_num_ops": [_P],
    "mb_graph_op_name": [_P, _I64], "mb_exec_create": [_P, _P, ctypes.c_char_p, _PP], "mb_exec_free": [_P],
    "mb_exec_set_transport": [_P, _P, _P], "mb_exec_bind_tensor": [_P, ctypes.c_char_p, _P], "mb_exec_run": [_P], "mb_exec_set_keep_values": [_P, _I],
    "mb_exec_executed": [_P], "mb_exec_bytes_sent": [_P], "mb_exec_seconds": [_P, _I],
    "mb_exec_value": [_P, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int), _PP, _P, ctypes.c_size_t,
                      ctypes.POINTER(ctypes.c_size_t), _I64P, ctypes.POINTER(ctypes.c_int)],
    "mb_exec_output": [_P, _I, ctypes.c_char_p, _PP], "mb_exec_output_key": [_P, _I, _I64],
    "mb_rep_index_axis0": [_P, _I64, _P, _PP], "mb_rep_shl_dim": [_P, _U64, _U64, _P, _PP],
}
_RESTYPES = {
    "mb_last_error": ctypes.c_char_p, "mb_ctx_stream": _P, "mb_ctx_launch_count": _U64,
    "mb_tensor_shape": _I64P, "mb_tensor_numel": _I64, "mb_tensor_nbytes": ctypes.c_size_t,
    "mb_tensor_device_ptr": _P, "mb_rep_share_tensor": _P,
    "mb_nccl_transport_fn": _P, "mb_peer_transport_fn": _P, "mb_party_bytes_sent": _U64, "mb_pshare_tensor": _P, "mb_party_stat": ctypes.c_double,
    "mb_graph_num_ops": _I64, "mb_graph_op_name": ctypes.c_char_p, "mb_exec_executed": _U64, "mb_exec_bytes_sent": _U64,
    "mb_exec_output_key": ctypes.c_char_p, "mb_exec_seconds": ctypes.c_double,
}


def _declare(l: ctypes.CDLL) -> None:
    for name, args in ABI.items():
        f = getattr(l, name)
        f.argtypes = args
        f.restype = _RESTYPES.get(name, _I)


def _shape_arr(shape: Sequence[int]):
    shape = tuple(int(s) for s in shape)
    return (ctypes.c_int64 * max(len(shape), 1))(*shape), len(shape)


class Tensor:
    """Device tensor handle: HostRing64Tensor / HostRing128Tensor / HostBitTensor / f64."""

    def __init__(self, sess: "Session", handle: int, owned: bool = True):
        self.sess = sess
        self.h = handle
        self.owned = owned

    def __del__(self):
        try:
            if self.owned and self.h and self.sess.ctx:
                lib().mb_tensor_free(self.sess.ctx, self.h)
        except Exception:
            pass
        self.h = None

    @property
    def dtype(self) -> int:
        return lib().mb_tensor_dtype(self.h)

    @property
    def shape(self) -> Tuple[int, ...]:
        r = lib().mb_tensor_rank(self.h)
        s = lib().mb_tensor_shape(self.h)
        return tuple(int(s[i]) for i in range(r))

    @property
    def numel(self) -> int:
        return int(lib().mb_tensor_numel(self.h))

    @property
    def nbytes(self) -> int:
        return int(lib().mb_tensor_nbytes(self.h))

    @property
    def device_ptr(self) -> int:
        return int(lib().mb_tensor_device_ptr(self.h) or 0)

    def to_host(self) -> np.ndarray:
        """ring64 -> uint64[shape]; ring128 -> uint64[shape+(2,)] = [lo, hi]; bit -> packed uint8;
        float64 -> float64[shape]."""
        dt, shape = self.dtype, self.shape
        if dt == RING64:
            out = np.empty(shape, dtype=np.uint64)
        elif dt == RING128:
            out = np.empty(shape + (2,), dtype=np.uint64)
        elif dt == FLOAT64:
            out = np.empty(shape, dtype=np.float64)
        else:
            out = np.empty(self.nbytes, dtype=np.uint8)
        self.sess._check(lib().mb_tensor_to_host(self.sess.ctx, self.h, out.ctypes.data_as(_P), out.nbytes))
        return out

    def to_ints(self) -> np.ndarray:
        a = self.to_host()
        if self.dtype == RING64:
            return np.array([int(v) for v in a.reshape(-1)], dtype=object).reshape(self.shape)
        if self.dtype == RING128:
            d = a.reshape(-1, 2)
            return np.array([int(lo) | (int(hi) << 64) for lo, hi in d], dtype=object).reshape(self.shape)
        raise TypeError("to_ints: ring tensors only")

    def bits(self) -> np.ndarray:
        """Unpacked 0/1 array of a bit tensor."""
        p = self.to_host()
        n = self.numel
        return np.unpackbits(p, bitorder="little")[:n].reshape(self.shape)


class RepTensor:
    """RepTensor (replicated/mod.rs:75-77) with the three distinct shares x0, x1, x2."""

    def __init__(self, sess: "Session", handle: int):
        self.sess = sess
        self.h = handle

    def __del__(self):
        try:
            if self.h and self.sess.ctx:
                lib().mb_rep_free(self.sess.ctx, self.h)
        except Exception:
            pass
        self.h = None

    def share(self, i: int) -> Tensor:
        self.sess._check(lib().mb_rep_materialize(self.sess.ctx, self.h))   # no-op unless the sharing was deferred
        return Tensor(self.sess, lib().mb_rep_share_tensor(self.h, i), owned=False)

    @property
    def shape(self):
        return self.share(0).shape

    @property
    def dtype(self):
        return self.share(0).dtype

    def reference_layout(self) -> List[List[np.ndarray]]:
        """[[x00,x10],[x11,x21],[x22,x02]] as the reference stores it."""
        s = [self.share(i).to_host() for i in range(3)]
        return [[s[0], s[1]], [s[1], s[2]], [s[2], s[0]]]


class Session:
    """One (device, session id, entropy source): the `sess` argument of the reference's kernels.

    `session_id` follows SessionId::try_from(&str) (computation.rs:108-124).  With
    `entropy_key` the PRF keys / unseeded sample seeds / sync keys are reproducible
    (determinism contract, include/moose_b200.h); without it they come from the OS.
    """

    def __init__(self, device: int = 0, session_id: str = "b200-bench", entropy_key: Optional[bytes] = None):
        self.ctx = None
        self._pinned = []
        ctx = ctypes.c_void_p()
        rc = lib().mb_ctx_create(device, ctypes.byref(ctx))
        if rc != 0:
            raise MooseB200Error(rc, (lib().mb_last_error(None) or b"").decode())
        self.ctx = ctx
        self._check(lib().mb_ctx_set_session(self.ctx, session_id.encode()))
        if entropy_key is not None:
            self._check(lib().mb_ctx_set_entropy_key(self.ctx, bytes(entropy_key)))

    def close(self):
        if self.ctx:
            for child in list(getattr(self, "_children", ())):  # native executors hold tensors of this context
                child._free()
            for p in self._pinned:
                lib().mb_host_free_pinned(self.ctx, p)
            self._pinned = []
            lib().mb_ctx_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc != 0:
            raise MooseB200Error(rc, (lib().mb_last_error(self.ctx) or b"").decode())

    # -- plumbing -------------------------------------------------------------
    def sync(self):
        self._check(lib().mb_ctx_sync(self.ctx))

    def timer_start(self):
        self._check(lib().mb_ctx_timer_start(self.ctx))

    def timer_stop(self) -> float:
        ms = ctypes.c_float()
        self._check(lib().mb_ctx_timer_stop(self.ctx, ctypes.byref(ms)))
        return float(ms.value)

    @property
    def launch_count(self) -> int:
        return int(lib().mb_ctx_launch_count(self.ctx))

    @property
    def stream(self) -> int:
        return int(lib().mb_ctx_stream(self.ctx) or 0)

    def profile_enable(self, on: bool = True):
        self._check(lib().mb_ctx_profile_enable(self.ctx, 1 if on else 0))

    def profile_read(self, kernel: int = 0) -> Tuple[float, int]:
        """(summed device ms, launches) of the ring-GEMM main kernel since the last read."""
        ms, n = ctypes.c_float(), ctypes.c_uint64()
        self._check(lib().mb_ctx_profile_read(self.ctx, kernel, ctypes.byref(ms), ctypes.byref(n)))
        return float(ms.value), int(n.value)

    def reserve_sms(self, count: int):
        self._check(lib().mb_ctx_reserve_sms(self.ctx, int(count)))

    def trace_enable(self, on: bool = True):
        self._check(lib().mb_ctx_trace_enable(self.ctx, 1 if on else 0))

    def trace_read(self) -> dict:
        """{launcher name: {"n": launches, "ms": summed device ms, "bytes": algorithmic HBM bytes, "aes_blocks": ...,
        "tensor_ops": ...}} since the trace was enabled / last read (work figures are 0 where a launcher states none)."""
        need = ctypes.c_size_t()
        buf = ctypes.create_string_buffer(1 << 16)
        self._check(lib().mb_ctx_trace_read(self.ctx, buf, len(buf), ctypes.byref(need)))
        out = {}
        for line in buf.value.decode().splitlines():
            name, n, ms, by, bl, ops = line.split("\t")
            out[name] = {"n": int(n), "ms": float(ms), "bytes": float(by), "aes_blocks": float(bl), "tensor_ops": float(ops)}
        return out

    def pinned_empty(self, shape, dtype) -> np.ndarray:
        """numpy array backed by page-locked host memory (freed with the session)."""
        dt = np.dtype(dtype)
        n = int(np.prod(shape, dtype=np.int64)) * dt.itemsize
        ptr = ctypes.c_void_p()
        self._check(lib().mb_host_alloc_pinned(self.ctx, n, ctypes.byref(ptr)))
        self._pinned.append(ptr)
        buf = (ctypes.c_uint8 * max(n, 1)).from_address(ptr.value)
        return np.frombuffer(buf, dtype=dt, count=n // dt.itemsize).reshape(shape)

    def set_fused_bits(self, on: bool):
        self._check(lib().mb_ctx_set_fused_bits(self.ctx, 1 if on else 0))

    def set_gemm_mode(self, mode: int):
        self._check(lib().mb_ctx_set_gemm_mode(self.ctx, mode))

    def session_id_raw(self) -> bytes:
        buf = ctypes.create_string_buffer(16)
        self._check(lib().mb_ctx_get_session_raw(self.ctx, buf))
        return buf.raw

    def _out(self):
        return ctypes.c_void_p()

    def _t(self, rc, out) -> Tensor:
        self._check(rc)
        return Tensor(self, out.value)

    def _r(self, rc, out) -> RepTensor:
        self._check(rc)
        return RepTensor(self, out.value)

    # -- tensors --------------------------------------------------------------
    def from_host(self, dtype: int, data: np.ndarray, shape: Optional[Sequence[int]] = None) -> Tensor:
        """data: uint64[shape] (ring64), uint64[shape+(2,)] (ring128), packed uint8 + shape (bit),
        float64[shape]."""
        in_shape = np.shape(data)  # np.ascontiguousarray promotes 0-d arrays to 1-d
        if dtype == RING64:
            a = np.ascontiguousarray(data, dtype=np.uint64)
            shape = in_shape if shape is None else tuple(shape)
        elif dtype == RING128:
            a = np.ascontiguousarray(data, dtype=np.uint64)
            shape = in_shape[:-1] if shape is None else tuple(shape)
        elif dtype == FLOAT64:
            a = np.ascontiguousarray(data, dtype=np.float64)
            shape = in_shape if shape is None else tuple(shape)
        else:
            a = np.ascontiguousarray(data, dtype=np.uint8)
            assert shape is not None, "bit tensors need an explicit shape"
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._t(lib().mb_tensor_from_host(self.ctx, dtype, rank, arr, a.ctypes.data_as(_P), ctypes.byref(out)), out)

    def from_ints(self, dtype: int, values, shape=None) -> Tensor:
        arr = np.array(values, dtype=object)
        shape = arr.shape if shape is None else tuple(shape)
        bits = 64 if dtype == RING64 else 128
        flat = [int(v) % (1 << bits) for v in arr.reshape(-1)]
        if dtype == RING64:
            return self.from_host(RING64, np.array(flat, dtype=np.uint64).reshape(shape))
        lo = np.array([v & ((1 << 64) - 1) for v in flat], dtype=np.uint64)
        hi = np.array([v >> 64 for v in flat], dtype=np.uint64)
        return self.from_host(RING128, np.stack([lo, hi], axis=-1).reshape(tuple(shape) + (2,)))

    def from_host_async(self, dtype: int, pinned: np.ndarray, shape: Optional[Sequence[int]] = None) -> Tensor:
        """H2D copy of a page-locked array (Session.pinned_empty) enqueued on this session's stream, no wait."""
        if shape is None:
            shape = pinned.shape[:-1] if dtype == RING128 else pinned.shape
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._t(lib().mb_tensor_from_host_async(self.ctx, dtype, rank, arr, pinned.ctypes.data_as(_P), ctypes.byref(out)), out)

    def to_host_async(self, x: Tensor, pinned_out: np.ndarray):
        self._check(lib().mb_tensor_to_host_async(self.ctx, x.h, pinned_out.ctypes.data_as(_P), pinned_out.nbytes))

    def wait(self, other: "Session"):
        """Work enqueued on this session from now on starts after everything enqueued on `other` so far."""
        self._check(lib().mb_ctx_wait(self.ctx, other.ctx))

    def from_bits(self, values) -> Tensor:
        a = np.asarray(values, dtype=np.uint8)
        return self.from_host(BIT, np.packbits(a.reshape(-1) & 1, bitorder="little"), a.shape)

    # -- HostPlacement kernels --------------------------------------------------
    def gen_key(self) -> bytes:
        buf = ctypes.create_string_buffer(16)
        self._check(lib().mb_prf_key_gen(self.ctx, buf))
        return buf.raw

    def sync_key_random(self) -> bytes:
        buf = ctypes.create_string_buffer(16)
        self._check(lib().mb_sync_key_random(self.ctx, buf))
        return buf.raw

    def derive_seed(self, sync_key: bytes, key: bytes) -> bytes:
        buf = ctypes.create_string_buffer(16)
        self._check(lib().mb_derive_seed(self.ctx, bytes(sync_key), bytes(key), buf))
        return buf.raw

    def sample_uniform_seeded(self, dtype: int, shape, seed: bytes, max_value: Optional[int] = None) -> Tensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        mv = -1 if max_value is None else int(max_value)
        return self._t(lib().mb_sample_seeded(self.ctx, dtype, rank, arr, bytes(seed), mv, ctypes.byref(out)), out)

    def sample_uniform_seeded_at(self, dtype: int, shape, seed: bytes, first_block: int) -> Tensor:
        """Uniform seeded stream entered at AES block `first_block` (mb_sample_seeded_at)."""
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._t(lib().mb_sample_seeded_at(self.ctx, dtype, rank, arr, bytes(seed), int(first_block), ctypes.byref(out)), out)

    def sample_bits_seeded(self, dtype: int, shape, seed: bytes) -> Tensor:
        return self.sample_uniform_seeded(dtype, shape, seed, max_value=1)

    def sample_uniform(self, dtype: int, shape, max_value: Optional[int] = None) -> Tensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        mv = -1 if max_value is None else int(max_value)
        return self._t(lib().mb_sample(self.ctx, dtype, rank, arr, mv, ctypes.byref(out)), out)

    def fill(self, dtype: int, value: int, shape) -> Tensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        value = int(value) % (1 << 128)
        return self._t(lib().mb_fill(self.ctx, dtype, rank, arr, value & ((1 << 64) - 1), value >> 64,
                                     ctypes.byref(out)), out)

    def _bin(self, fn, x: Tensor, y: Tensor) -> Tensor:
        out = self._out()
        return self._t(fn(self.ctx, x.h, y.h, ctypes.byref(out)), out)

    def add(self, x, y): return self._bin(lib().mb_add, x, y)
    def sub(self, x, y): return self._bin(lib().mb_sub, x, y)
    def mul(self, x, y): return self._bin(lib().mb_mul, x, y)
    def dot(self, x, y): return self._bin(lib().mb_dot, x, y)
    def xor(self, x, y): return self._bin(lib().mb_xor, x, y)
    def and_(self, x, y): return self._bin(lib().mb_and, x, y)
    def or_(self, x, y): return self._bin(lib().mb_or, x, y)

    def dot_noalloc(self, x: Tensor, y: Tensor, out: Tensor) -> None:
        """C = x . y into the caller-provided `out` through mb_dot_noalloc (no allocation inside the call)."""
        M, K = x.shape
        N = y.shape[1]
        self._check(lib().mb_dot_noalloc(self.ctx, x.dtype, x.device_ptr, y.device_ptr, out.device_ptr, M, K, N))

    def alloc(self, dtype: int, shape) -> Tensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._t(lib().mb_tensor_alloc(self.ctx, dtype, rank, arr, ctypes.byref(out)), out)

    def less(self, x, y): return self._bin(lib().mb_less, x, y)
    def greater(self, x, y): return self._bin(lib().mb_greater, x, y)

    def host_mux(self, s: Tensor, x: Tensor, y: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_mux(self.ctx, s.h, x.h, y.h, ctypes.byref(out)), out)

    def sign(self, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_sign(self.ctx, x.h, ctypes.byref(out)), out)

    def cast_ring128_to_ring64(self, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_cast_ring128_to_ring64(self.ctx, x.h, ctypes.byref(out)), out)

    def host_maximum(self, xs: Sequence[Tensor]) -> Tensor:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._t(lib().mb_maximum(self.ctx, arr, len(xs), ctypes.byref(out)), out)

    def neg(self, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_neg(self.ctx, x.h, ctypes.byref(out)), out)

    def shl(self, amount: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_shl(self.ctx, amount, x.h, ctypes.byref(out)), out)

    def shr(self, amount: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_shr(self.ctx, amount, x.h, ctypes.byref(out)), out)

    def add_n(self, xs: Sequence[Tensor]) -> Tensor:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._t(lib().mb_add_n(self.ctx, arr, len(xs), ctypes.byref(out)), out)

    def sum(self, x: Tensor, axis: Optional[int] = None) -> Tensor:
        out = self._out()
        return self._t(lib().mb_sum(self.ctx, x.h, -1 if axis is None else axis, ctypes.byref(out)), out)

    def broadcast(self, x: Tensor, shape) -> Tensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._t(lib().mb_broadcast(self.ctx, x.h, rank, arr, ctypes.byref(out)), out)

    def reshape(self, x: Tensor, shape) -> Tensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._t(lib().mb_reshape(self.ctx, x.h, rank, arr, ctypes.byref(out)), out)

    def fixedpoint_encode(self, ring: int, frac: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_fixedpoint_encode(self.ctx, ring, frac, x.h, ctypes.byref(out)), out)

    def fixedpoint_decode(self, frac: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_fixedpoint_decode(self.ctx, frac, x.h, ctypes.byref(out)), out)

    def bit_decompose(self, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_bit_decompose(self.ctx, x.h, ctypes.byref(out)), out)

    def bit_extract(self, bit_idx: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_bit_extract(self.ctx, bit_idx, x.h, ctypes.byref(out)), out)

    def ring_inject(self, ring: int, bit_idx: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_ring_inject(self.ctx, ring, bit_idx, x.h, ctypes.byref(out)), out)

    def shl_dim(self, amount: int, bit_length: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_shl_dim(self.ctx, amount, bit_length, x.h, ctypes.byref(out)), out)

    def index_axis0(self, index: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_index_axis0(self.ctx, index, x.h, ctypes.byref(out)), out)

    def index_axis(self, axis: int, index: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_index_axis(self.ctx, axis, index, x.h, ctypes.byref(out)), out)

    def expand_dims(self, axis, x: Tensor) -> Tensor:
        axes = [axis] if isinstance(axis, int) else list(axis)
        arr = (ctypes.c_int64 * max(len(axes), 1))(*axes)
        out = self._out()
        return self._t(lib().mb_expand_dims(self.ctx, arr, len(axes), x.h, ctypes.byref(out)), out)

    def squeeze(self, axis: Optional[int], x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_squeeze(self.ctx, -1 if axis is None else axis, x.h, ctypes.byref(out)), out)

    def transpose(self, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_transpose(self.ctx, x.h, ctypes.byref(out)), out)

    def concat(self, axis: int, xs: Sequence[Tensor]) -> Tensor:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._t(lib().mb_concat(self.ctx, axis, arr, len(xs), ctypes.byref(out)), out)

    def slice(self, info, x: Tensor) -> Tensor:
        """`info`: one (start, end | None, step | None) triple or python slice per axis (SliceInfo, host/mod.rs:153)."""
        trip = [(e.start or 0, e.stop, e.step) if isinstance(e, slice) else tuple(e) for e in info]
        n = len(trip)
        start = (ctypes.c_int64 * max(n, 1))(*[t[0] for t in trip])
        end = (ctypes.c_int64 * max(n, 1))(*[0 if t[1] is None else t[1] for t in trip])
        has_end = bytes(0 if t[1] is None else 1 for t in trip) + b"\0"
        step = (ctypes.c_int64 * max(n, 1))(*[1 if t[2] is None else t[2] for t in trip])
        out = self._out()
        return self._t(lib().mb_slice(self.ctx, n, start, end, has_end, step, x.h, ctypes.byref(out)), out)

    def diag(self, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_diag(self.ctx, x.h, ctypes.byref(out)), out)

    def select(self, axis: int, index_bits: Tensor, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_select(self.ctx, axis, index_bits.h, x.h, ctypes.byref(out)), out)

    def div(self, x: Tensor, y: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_div(self.ctx, x.h, y.h, ctypes.byref(out)), out)

    def bit_decompose_ring(self, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_bit_decompose_ring(self.ctx, x.h, ctypes.byref(out)), out)

    def fixedpoint_mean(self, axis: Optional[int], scaling_base: int, scaling_exp: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_fixedpoint_mean(self.ctx, -1 if axis is None else axis, scaling_base, scaling_exp,
                                                x.h, ctypes.byref(out)), out)

    def fixedpoint_argmax(self, axis: int, x: Tensor) -> Tensor:
        out = self._out()
        return self._t(lib().mb_fixedpoint_argmax(self.ctx, axis, x.h, ctypes.byref(out)), out)

    # -- storage (LoadOp / SaveOp) ------------------------------------------------
    def load_npy(self, path: str, ring: Optional[int] = None, frac: int = 0) -> Tensor:
        """float64 tensor, or (ring given) the fixed-point encoding computed while the file streams in."""
        out = self._out()
        if ring is None:
            return self._t(lib().mb_load_npy(self.ctx, path.encode(), ctypes.byref(out)), out)
        return self._t(lib().mb_load_npy_fixed(self.ctx, path.encode(), ring, frac, ctypes.byref(out)), out)

    def save_npy(self, x: Tensor, path: str):
        self._check(lib().mb_save_npy(self.ctx, x.h, path.encode()))

    def load_csv(self, path: str, columns: Sequence[str] = ()) -> Tensor:
        arr = (ctypes.c_char_p * max(len(columns), 1))(*[c.encode() for c in columns])
        out = self._out()
        return self._t(lib().mb_load_csv(self.ctx, path.encode(), ctypes.cast(arr, _PP), len(columns), ctypes.byref(out)), out)

    def save_csv(self, x: Tensor, path: str):
        self._check(lib().mb_save_csv(self.ctx, x.h, path.encode()))

    # -- ReplicatedPlacement kernels ---------------------------------------------
    def rep_setup(self):
        self._check(lib().mb_rep_setup(self.ctx))

    def rep_from_shares(self, x0: Tensor, x1: Tensor, x2: Tensor) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_from_shares(self.ctx, x0.h, x1.h, x2.h, ctypes.byref(out)), out)

    def share(self, x: Tensor, owner: int = 0, deferred: bool = False) -> RepTensor:
        """ShareOp.  deferred=True (placement members only): the key is drawn now, the shares are expanded inside the
        skinny dot that consumes them (rep_dot_concat / fixed_dot_concat) or on first access."""
        out = self._out()
        if deferred:
            return self._r(lib().mb_rep_share_deferred(self.ctx, x.h, owner, ctypes.byref(out)), out)
        return self._r(lib().mb_rep_share(self.ctx, x.h, owner, ctypes.byref(out)), out)

    def reveal(self, x: RepTensor, receiver: int = 0) -> Tensor:
        out = self._out()
        return self._t(lib().mb_rep_reveal(self.ctx, x.h, receiver, ctypes.byref(out)), out)

    def _rbin(self, fn, x, y) -> RepTensor:
        out = self._out()
        return self._r(fn(self.ctx, x.h, y.h, ctypes.byref(out)), out)

    def rep_add(self, x, y): return self._rbin(lib().mb_rep_add, x, y)
    def rep_sub(self, x, y): return self._rbin(lib().mb_rep_sub, x, y)
    def rep_mul(self, x, y): return self._rbin(lib().mb_rep_mul, x, y)
    def rep_dot(self, x, y): return self._rbin(lib().mb_rep_dot, x, y)
    def rep_less(self, x, y): return self._rbin(lib().mb_rep_less, x, y)
    def rep_add_public(self, x, pub): return self._rbin(lib().mb_rep_add_public, x, pub)
    def rep_sub_public(self, x, pub): return self._rbin(lib().mb_rep_sub_public, x, pub)
    def rep_mul_public(self, x, pub): return self._rbin(lib().mb_rep_mul_public, x, pub)

    def rep_public_sub(self, pub: Tensor, x: RepTensor) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_public_sub(self.ctx, pub.h, x.h, ctypes.byref(out)), out)

    def rep_neg(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_neg(self.ctx, x.h, ctypes.byref(out)), out)

    def rep_shl(self, amount: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_shl(self.ctx, amount, x.h, ctypes.byref(out)), out)

    def trunc_pr(self, amount: int, x: RepTensor) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_trunc_pr(self.ctx, amount, x.h, ctypes.byref(out)), out)

    def rep_bit_decompose(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_bit_decompose(self.ctx, x.h, ctypes.byref(out)), out)

    def msb(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_msb(self.ctx, x.h, ctypes.byref(out)), out)

    def rep_ring_inject(self, ring: int, bit_idx: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_ring_inject(self.ctx, ring, bit_idx, x.h, ctypes.byref(out)), out)

    def mux(self, s, x, y) -> RepTensor:
        """s*(x-y)+y; `s` is a replicated ring value or replicated bits (control_flow.rs:9-45)."""
        out = self._out()
        return self._r(lib().mb_rep_mux(self.ctx, s.h, x.h, y.h, ctypes.byref(out)), out)

    def bit_compose(self, ring: int, x_bits) -> RepTensor:
        """BitComposeOp::rep_kernel (replicated/bits.rs:23-46): planes [k, shape...] -> ring value."""
        out = self._out()
        return self._r(lib().mb_rep_bit_compose(self.ctx, ring, x_bits.h, ctypes.byref(out)), out)

    def maximum(self, xs: Sequence[RepTensor]) -> RepTensor:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._r(lib().mb_rep_maximum(self.ctx, arr, len(xs), ctypes.byref(out)), out)

    def relu(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_relu(self.ctx, x.h, ctypes.byref(out)), out)

    def rep_fill(self, dtype: int, value: int, shape) -> RepTensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        value = int(value) % (1 << 128)
        return self._r(lib().mb_rep_fill(self.ctx, dtype, rank, arr, value & ((1 << 64) - 1), value >> 64,
                                         ctypes.byref(out)), out)

    def rep_greater(self, x, y): return self._rbin(lib().mb_rep_greater, x, y)

    def rep_div(self, i_prec: int, f_prec: int, x, y) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_div(self.ctx, i_prec, f_prec, x.h, y.h, ctypes.byref(out)), out)

    def _fixed_unary(self, fn, i_prec, f_prec, x) -> RepTensor:
        out = self._out()
        return self._r(fn(self.ctx, i_prec, f_prec, x.h, ctypes.byref(out)), out)

    def rep_pow2(self, i_prec, f_prec, x): return self._fixed_unary(lib().mb_rep_pow2, i_prec, f_prec, x)
    def rep_exp(self, i_prec, f_prec, x): return self._fixed_unary(lib().mb_rep_exp, i_prec, f_prec, x)
    def sigmoid(self, i_prec, f_prec, x): return self._fixed_unary(lib().mb_rep_sigmoid, i_prec, f_prec, x)
    def rep_log2(self, i_prec, f_prec, x): return self._fixed_unary(lib().mb_rep_log2, i_prec, f_prec, x)
    def rep_log(self, i_prec, f_prec, x): return self._fixed_unary(lib().mb_rep_log, i_prec, f_prec, x)
    def rep_sqrt(self, i_prec, f_prec, x): return self._fixed_unary(lib().mb_rep_sqrt, i_prec, f_prec, x)
    def rep_equal(self, x, y): return self._rbin(lib().mb_rep_equal, x, y)

    # share-wise shape / reduction operators (replicated/ops.rs, arith.rs:123-195)
    def rep_index_axis(self, axis: int, index: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_index_axis(self.ctx, axis, index, x.h, ctypes.byref(out)), out)

    def rep_expand_dims(self, axis, x) -> RepTensor:
        axes = [axis] if isinstance(axis, int) else list(axis)
        arr = (ctypes.c_int64 * max(len(axes), 1))(*axes)
        out = self._out()
        return self._r(lib().mb_rep_expand_dims(self.ctx, arr, len(axes), x.h, ctypes.byref(out)), out)

    def rep_squeeze(self, axis: Optional[int], x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_squeeze(self.ctx, -1 if axis is None else axis, x.h, ctypes.byref(out)), out)

    def rep_transpose(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_transpose(self.ctx, x.h, ctypes.byref(out)), out)

    def rep_reshape(self, x, shape) -> RepTensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._r(lib().mb_rep_reshape(self.ctx, x.h, rank, arr, ctypes.byref(out)), out)

    def rep_broadcast(self, x, shape) -> RepTensor:
        arr, rank = _shape_arr(shape)
        out = self._out()
        return self._r(lib().mb_rep_broadcast(self.ctx, x.h, rank, arr, ctypes.byref(out)), out)

    def rep_slice(self, info, x) -> RepTensor:
        trip = [(e.start or 0, e.stop, e.step) if isinstance(e, slice) else tuple(e) for e in info]
        n = len(trip)
        start = (ctypes.c_int64 * max(n, 1))(*[t[0] for t in trip])
        end = (ctypes.c_int64 * max(n, 1))(*[0 if t[1] is None else t[1] for t in trip])
        has_end = bytes(0 if t[1] is None else 1 for t in trip) + b"\0"
        step = (ctypes.c_int64 * max(n, 1))(*[1 if t[2] is None else t[2] for t in trip])
        out = self._out()
        return self._r(lib().mb_rep_slice(self.ctx, n, start, end, has_end, step, x.h, ctypes.byref(out)), out)

    def rep_diag(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_diag(self.ctx, x.h, ctypes.byref(out)), out)

    def rep_concat(self, axis: int, xs) -> RepTensor:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._r(lib().mb_rep_concat(self.ctx, axis, arr, len(xs), ctypes.byref(out)), out)

    def rep_dot_concat(self, xs, y) -> RepTensor:
        """dot(concatenate(xs, axis=1), y); the concatenation is not written when y is skinny."""
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._r(lib().mb_rep_dot_concat(self.ctx, arr, len(xs), y.h, ctypes.byref(out)), out)

    def fixed_dot(self, amount: int, x, y) -> RepTensor:
        """trunc_pr(amount, rep_dot(x, y)) in one call (same shares; the product's zero share is never expanded)."""
        out = self._out()
        return self._r(lib().mb_rep_fixed_dot(self.ctx, amount, x.h, y.h, ctypes.byref(out)), out)

    def fixed_dot_concat(self, amount: int, xs, y) -> RepTensor:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._r(lib().mb_rep_fixed_dot_concat(self.ctx, amount, arr, len(xs), y.h, ctypes.byref(out)), out)

    def fixed_mul(self, amount: int, x, y) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_fixed_mul(self.ctx, amount, x.h, y.h, ctypes.byref(out)), out)

    def rep_add_n(self, xs) -> RepTensor:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = self._out()
        return self._r(lib().mb_rep_add_n(self.ctx, arr, len(xs), ctypes.byref(out)), out)

    def rep_sum(self, x, axis: Optional[int] = None) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_sum(self.ctx, x.h, -1 if axis is None else axis, ctypes.byref(out)), out)

    def rep_cast_ring128_to_ring64(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_cast_ring128_to_ring64(self.ctx, x.h, ctypes.byref(out)), out)

    def rep_fixedpoint_mean(self, axis: Optional[int], scaling_base: int, scaling_exp: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_fixedpoint_mean(self.ctx, -1 if axis is None else axis, scaling_base, scaling_exp,
                                                    x.h, ctypes.byref(out)), out)

    def rep_abs(self, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_abs(self.ctx, x.h, ctypes.byref(out)), out)

    def softmax(self, axis: int, upmost_index: int, i_prec: int, f_prec: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_softmax(self.ctx, axis, upmost_index, i_prec, f_prec, x.h, ctypes.byref(out)), out)

    def argmax(self, axis: int, upmost_index: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_argmax(self.ctx, axis, upmost_index, x.h, ctypes.byref(out)), out)

    def rep_index_axis0(self, index: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_index_axis0(self.ctx, index, x.h, ctypes.byref(out)), out)

    def rep_shl_dim(self, amount: int, bit_length: int, x) -> RepTensor:
        out = self._out()
        return self._r(lib().mb_rep_shl_dim(self.ctx, amount, bit_length, x.h, ctypes.byref(out)), out)
