"""Replay of lowered (host-dialect) Moose computations: Python face of the C++ reader + executor in
libmoose_b200 (csrc/executor.cpp, entry points mb_graph_* / mb_exec_* of include/moose_b200.h).

The executor stands where the reference's executors stand (moose/src/execution/synchronous.rs:347-361 and
asynchronous.rs:453-531): it walks a parsed computation in topological order and dispatches every operation to the
host-kernel entry point that replaces its Rust kernel -- the same (operator, placement, operand types) -> kernel
mapping as `DispatchKernel::compile` (moose/src/kernels/mod.rs:34-106).  Parsing, scheduling, dispatch and the value
store are native; this module only marshals graphs / tensors and binds the transport.

Two modes:
  * all parties on one GPU: every `@Host(role)` op runs on the one session; a Send/Receive pair
    (moose/src/networking/mod.rs:16-54) is a hand-off keyed by the rendezvous key, like the reference's in-memory
    LocalSyncNetworking (networking/local.rs:10-53);
  * party per GPU: each process owns one role and one GPU, runs only that role's ops, and Send/Receive become NCCL
    send/recv (torch.distributed) between the parties' GPUs, straight from / into the tensors' device memory.
    Receives directly follow their sends in the shared topological order, so all parties post matching transfers
    in the same order.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Union

import numpy as np

from . import Session, Tensor, lib
from .textual import Operation, dumps_text


class ExecutionError(RuntimeError):
    pass


def toposort(ops: List[Operation]) -> List[Operation]:
    """Stable topological order; every Receive is scheduled right after its Send (same rendezvous key)."""
    by_name = {op.name: op for op in ops}
    send_of = {op.attrs["rendezvous_key"]: op.name for op in ops if op.kind == "Send"}
    deps = {}
    for op in ops:
        d = [i for i in op.inputs if i in by_name]
        if op.kind == "Receive" and op.attrs["rendezvous_key"] in send_of:
            d.append(send_of[op.attrs["rendezvous_key"]])
        deps[op.name] = d
    users: Dict[str, List[str]] = {op.name: [] for op in ops}
    indeg = {}
    for op in ops:
        indeg[op.name] = len(deps[op.name])
        for d in deps[op.name]:
            users[d].append(op.name)
    pos = {op.name: i for i, op in enumerate(ops)}
    import heapq
    ready = [(pos[n], n) for n, k in indeg.items() if k == 0]
    heapq.heapify(ready)
    order = []
    while ready:
        _, n = heapq.heappop(ready)
        op = by_name[n]
        order.append(op)
        # a Send releases its Receive immediately (priority -1 keeps the pair adjacent)
        for u in users[n]:
            indeg[u] -= 1
            if indeg[u] == 0:
                pri = -1 if (op.kind == "Send" and by_name[u].kind == "Receive") else pos[u]
                heapq.heappush(ready, (pri, u))
    if len(order) != len(ops):
        raise ExecutionError("computation has a cycle or a dangling input")
    return order


class NcclComm:
    """Send/Receive between parties over NCCL (torch.distributed P2P).  Wire format: an int64[10] header
    (kind, dtype, rank, shape[6], nbytes) followed by the raw payload bytes."""

    K_TENSOR, K_BYTES, K_SHAPE, K_UNIT, K_STRING = 0, 1, 2, 3, 4

    def __init__(self, dist, sess: Optional[Session], role_to_rank: Dict[str, int], device):
        """`device`: CUDA ordinal for NCCL; "cpu" for the gloo tests of the wire format (keys, seeds, shapes,
        units -- tensors need a GPU session)."""
        import torch
        self.torch, self.dist, self.sess, self.role_to_rank = torch, dist, sess, role_to_rank
        self.dev = torch.device("cpu") if device == "cpu" else torch.device("cuda", device)
        self.bytes_sent = 0
        self._posted, self._completed = [], 0  # device sends in flight (Work handles) / completed so far

    def _hdr(self, vals):
        return self.torch.tensor(list(vals) + [0] * (10 - len(vals)), dtype=self.torch.int64, device=self.dev)

    def send(self, value, receiver: str):
        t, dst = self.torch, self.role_to_rank[receiver]
        if isinstance(value, Tensor):
            shape = list(value.shape)
            hdr = [self.K_TENSOR, value.dtype, len(shape)] + shape + [0] * (6 - len(shape)) + [value.nbytes]
            buf = t.empty(max(value.nbytes, 1), dtype=t.uint8, device=self.dev)
            self.sess._check(lib().mb_tensor_copy_to_device(self.sess.ctx, value.h, buf.data_ptr()))
            self.sess.sync()
            self.dist.send(self._hdr(hdr), dst)
            if value.nbytes:
                self.dist.send(buf, dst)
            self.bytes_sent += value.nbytes
        elif isinstance(value, (bytes, bytearray)):
            self.dist.send(self._hdr([self.K_BYTES, 0, 0, 0, 0, 0, 0, 0, 0, len(value)]), dst)
            self.dist.send(t.tensor(list(value), dtype=t.uint8, device=self.dev), dst)
            self.bytes_sent += len(value)
        elif isinstance(value, tuple):
            shape = list(value)
            self.dist.send(self._hdr([self.K_SHAPE, 0, len(shape)] + shape + [0] * (6 - len(shape)) + [0]), dst)
        elif value is None:
            self.dist.send(self._hdr([self.K_UNIT]), dst)
        else:
            raise ExecutionError(f"cannot send value of type {type(value)}")

    def transport(self, _user, mode, peer, buf, nbytes, on_device):
        """mb_transport_fn: raw buffers of the native executor.  Device payloads go straight from / into the
        tensor's memory; host buffers (headers, keys, seeds) are staged through a small device tensor for NCCL.
        Modes: 0 receive, 1 send, 2 poll completed device sends, 3 wait for all of them, 4 grouped send + receive,
        5 capabilities."""
        try:
            t = self.torch
            if mode == 5:  # capabilities: none (transfers run on NCCL's stream)
                ctypes.cast(buf, ctypes.POINTER(ctypes.c_int64))[0] = 0
                return 0
            if mode in (2, 3):
                while self._posted and (mode == 3 or self._posted[0].is_completed()):
                    self._posted.pop(0).wait()
                    if mode == 3 and self.dev.type == "cuda" and not self._posted:
                        t.cuda.synchronize(self.dev)
                    self._completed += 1
                ctypes.cast(buf, ctypes.POINTER(ctypes.c_int64))[0] = self._completed
                return 0
            if mode == 4:  # grouped send + receive of two device buffers (ring exchange of the party runtime)
                d = ctypes.cast(buf, ctypes.POINTER(SendRecv))[0]
                sview = t.as_tensor(_DeviceBytes(d.send_buf, d.send_bytes), device=self.dev)
                rview = t.as_tensor(_DeviceBytes(d.recv_buf, d.recv_bytes), device=self.dev)
                ops = [self.dist.P2POp(self.dist.isend, sview, self.role_to_rank[d.send_peer.decode()]),
                       self.dist.P2POp(self.dist.irecv, rview, self.role_to_rank[d.recv_peer.decode()])]
                for req in self.dist.batch_isend_irecv(ops):
                    req.wait()
                done = t.cuda.Event()
                done.record()
                done.synchronize()
                return 0
            other = self.role_to_rank[peer.decode()]
            if on_device:
                view = t.as_tensor(_DeviceBytes(buf, nbytes), device=self.dev)
                if mode == 1:
                    self._posted.append(self.dist.isend(view, other))  # asynchronous on the NCCL stream
                else:
                    # wait for THIS transfer only (not for the party's own kernels on the context's stream)
                    self.dist.irecv(view, other).wait()
                    done = t.cuda.Event()
                    done.record()
                    done.synchronize()
                return 0
            host = (ctypes.c_uint8 * nbytes).from_address(buf)
            if mode == 1:
                self.dist.send(t.frombuffer(bytearray(host), dtype=t.uint8).to(self.dev), other)
            else:
                staged = t.empty(nbytes, dtype=t.uint8, device=self.dev)
                self.dist.recv(staged, other)
                ctypes.memmove(buf, bytes(staged.cpu().numpy().tobytes()), nbytes)
            return 0
        except Exception as e:  # an exception must not cross the C boundary
            import sys
            print(f"transport error: {e!r}", file=sys.stderr, flush=True)
            return 1

    def flush(self):
        """All posted transfers have completed (their buffers may be freed / reused afterwards)."""
        if self.dev.type == "cuda":
            self.torch.cuda.synchronize(self.dev)

    def recv(self, sender: str):
        t, src = self.torch, self.role_to_rank[sender]
        hdr = t.zeros(10, dtype=t.int64, device=self.dev)
        self.dist.recv(hdr, src)
        h = [int(v) for v in hdr.cpu()]
        kind, dtype, rank, shape, nbytes = h[0], h[1], h[2], h[3:3 + h[2]], h[9]
        if kind == self.K_TENSOR:
            out = self.sess.alloc(dtype, shape)
            if nbytes:
                buf = t.empty(nbytes, dtype=t.uint8, device=self.dev)
                self.dist.recv(buf, src)
                if self.dev.type == "cuda":
                    t.cuda.synchronize(self.dev)
                self.sess._check(lib().mb_tensor_copy_from_device(self.sess.ctx, out.h, buf.data_ptr()))
                self.sess.sync()
            return out
        if kind == self.K_BYTES:
            buf = t.empty(nbytes, dtype=t.uint8, device=self.dev)
            self.dist.recv(buf, src)
            return bytes(buf.cpu().tolist())
        if kind == self.K_SHAPE:
            return tuple(shape)
        return None


class SendRecv(ctypes.Structure):
    """mb_sendrecv (include/moose_b200.h): argument of transport mode 4."""
    _fields_ = [("send_peer", ctypes.c_char_p), ("send_buf", ctypes.c_void_p), ("send_bytes", ctypes.c_size_t),
                ("recv_peer", ctypes.c_char_p), ("recv_buf", ctypes.c_void_p), ("recv_bytes", ctypes.c_size_t)]


TRANSPORT_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_void_p,
                                ctypes.c_size_t, ctypes.c_int)


class _DeviceBytes:
    """Zero-copy view of `nbytes` of device memory for torch (CUDA array interface)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class _Values:
    """Results of the executed operations by name (fetched from the native value store on demand)."""

    def __init__(self, ex: "Executor"):
        import weakref
        self._ex = weakref.ref(ex)  # no reference cycle: dropping the executor frees its device tensors at once

    def __getitem__(self, name: str):
        ex = self._ex()
        kind, rank, nbytes = ctypes.c_int(-1), ctypes.c_int(0), ctypes.c_size_t(0)
        handle = ctypes.c_void_p()
        buf = (ctypes.c_uint8 * 4096)()
        shape = (ctypes.c_int64 * 6)()
        ex.sess._check(lib().mb_exec_value(ex._exec, name.encode(), ctypes.byref(kind), ctypes.byref(handle), buf, 4096,
                                           ctypes.byref(nbytes), shape, ctypes.byref(rank)))
        if kind.value == 0:
            return Tensor(ex.sess, handle.value)
        if kind.value == 1:
            return bytes(buf[:nbytes.value])
        if kind.value == 2:
            return tuple(int(shape[i]) for i in range(rank.value))
        if kind.value == 4:
            return bytes(buf[:nbytes.value]).decode()
        if kind.value == 3:
            return None
        raise KeyError(f"{name} was not executed in this process")


class Executor:
    def __init__(self, sess: Session, role: Optional[str] = None, comm: Optional[NcclComm] = None,
                 overrides: Optional[Dict[str, Tensor]] = None, keep_values: bool = True):
        """`overrides` replaces the result of named ops (e.g. the Constant inputs of a compiled graph by larger
        tensors of the same rank: lowered graphs are shape-agnostic).  keep_values=False releases every intermediate
        value after its last use (only outputs / saved values can be read afterwards)."""
        self.sess, self.role, self.comm, self.keep_values = sess, role, comm, keep_values
        self.overrides = overrides or {}
        self.values = _Values(self)
        self.outputs: Dict[str, Tensor] = {}
        self.saved: Dict[str, Tensor] = {}
        self.executed = 0
        self._graph = ctypes.c_void_p()
        self._exec = ctypes.c_void_p()
        self._cb = None
        if not hasattr(sess, "_children"):
            import weakref
            sess._children = weakref.WeakSet()
        sess._children.add(self)

    def _free(self):
        if not self.sess.ctx:  # the context is gone (and with it the stream the tensors would be freed on)
            self._exec, self._graph = ctypes.c_void_p(), ctypes.c_void_p()
            return
        if self._exec:
            lib().mb_exec_free(self._exec)
            self._exec = ctypes.c_void_p()
        if self._graph:
            lib().mb_graph_free(self._graph)
            self._graph = ctypes.c_void_p()

    def close(self):
        self.outputs, self.saved = {}, {}
        self._free()

    def __del__(self):
        try:
            self._free()
        except Exception:
            pass

    def order(self) -> List[str]:
        """Operation names in the native execution order."""
        n = lib().mb_graph_num_ops(self._graph)
        return [lib().mb_graph_op_name(self._graph, i).decode() for i in range(n)]

    def load(self, ops: Union[str, List[Operation]]):
        self._free()
        text = ops if isinstance(ops, str) else dumps_text(ops)
        self.sess._check(lib().mb_graph_parse(self.sess.ctx, text.encode(), ctypes.byref(self._graph)))
        self.sess._check(lib().mb_exec_create(self.sess.ctx, self._graph, self.role.encode() if self.role else None,
                                              ctypes.byref(self._exec)))
        self.sess._check(lib().mb_exec_set_keep_values(self._exec, int(self.keep_values)))
        if self.comm is not None and hasattr(self.comm, "fn"):   # a transport of the library: C function + handle
            self._cb = ctypes.c_void_p(self.comm.fn)
            self.sess._check(lib().mb_exec_set_transport(self._exec, self._cb, self.comm.handle))
        elif self.comm is not None:
            self._cb = TRANSPORT_FN(self.comm.transport)
            self.sess._check(lib().mb_exec_set_transport(self._exec, self._cb, None))
        return self

    def run(self, ops: Union[str, List[Operation], None] = None, arguments: Optional[Dict[str, Tensor]] = None):
        if ops is not None:
            self.load(ops)
        for name, t in list(self.overrides.items()) + list((arguments or {}).items()):
            self.sess._check(lib().mb_exec_bind_tensor(self._exec, name.encode(), t.h))
        native = self.comm is not None and hasattr(self.comm, "fn")
        if self.comm is not None and not native:
            self.comm.flush()
            self.comm._posted, self.comm._completed = [], 0  # the executor counts completed sends per run
        try:
            self.sess._check(lib().mb_exec_run(self._exec))
        except Exception as e:
            raise ExecutionError(str(e)) from e
        self.executed = int(lib().mb_exec_executed(self._exec))
        self.transport_seconds = float(lib().mb_exec_seconds(self._exec, 0))
        self.sync_seconds = float(lib().mb_exec_seconds(self._exec, 1))
        if self.comm is not None and not native:
            self.comm.flush()
        if self.comm is not None:
            self.comm.bytes_sent = getattr(self.comm, "bytes_sent", 0) + int(lib().mb_exec_bytes_sent(self._exec))
        for saved, store in ((0, self.outputs), (1, self.saved)):
            i = 0
            while True:
                key = lib().mb_exec_output_key(self._exec, saved, i)
                if not key:
                    break
                h = ctypes.c_void_p()
                self.sess._check(lib().mb_exec_output(self._exec, saved, key, ctypes.byref(h)))
                store[key.decode()] = Tensor(self.sess, h.value)
                i += 1
        return self.outputs
