"""Replay executor for lowered (host-dialect) Moose computations over libmoose_b200.

Stand-in for the reference's executors (moose/src/execution/synchronous.rs:347-361 and
asynchronous.rs:453-531) on the other side of the C-ABI boundary: those are Rust and cannot be built in this
image, so this module walks a parsed computation (moose_b200/textual.py) in topological order and dispatches
every operation to the host-kernel entry point that replaces its Rust kernel -- the same (operator,
placement, operand types) -> kernel mapping as `DispatchKernel::compile` (moose/src/kernels/mod.rs:34-106).

Two modes:
  * all parties on one GPU: every `@Host(role)` op runs on the one session; a Send/Receive pair
    (moose/src/networking/mod.rs:16-54) is a handle hand-off keyed by the rendezvous key, exactly like the
    reference's in-memory LocalSyncNetworking (networking/local.rs:10-53);
  * party per GPU: each process owns one role and one GPU, runs only that role's ops, and maps Send/Receive
    to NCCL send/recv (torch.distributed) between the parties' GPUs.  Receives directly follow their sends in
    the shared topological order, so all parties post matching transfers in the same order.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np

from . import BIT, FLOAT64, RING64, RING128, Session, Tensor, lib, _P
from .textual import Operation, attr_hex16, attr_int, attr_int_list, attr_slice, attr_str, parse_constant

_RING_TYPES = {"HostRing64Tensor": RING64, "HostRing128Tensor": RING128, "Ring64Tensor": RING64, "Ring128Tensor": RING128}
_BIT_TYPES = {"HostBitTensor", "BitTensor"}


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


class Executor:
    def __init__(self, sess: Session, role: Optional[str] = None, comm: Optional[NcclComm] = None,
                 overrides: Optional[Dict[str, object]] = None):
        """`overrides` replaces the result of named ops (e.g. the Constant inputs of a compiled graph by larger
        tensors of the same rank: lowered graphs are shape-agnostic)."""
        self.sess, self.role, self.comm = sess, role, comm
        self.overrides = overrides or {}
        self.values: Dict[str, object] = {}
        self.outputs: Dict[str, object] = {}
        self.saved: Dict[str, object] = {}
        self.mailbox: Dict[str, object] = {}
        self.executed = 0

    # -- helpers -------------------------------------------------------------------------------
    def _dtype_of(self, ty: str) -> int:
        if ty in _RING_TYPES:
            return _RING_TYPES[ty]
        if ty in _BIT_TYPES:
            return BIT
        raise ExecutionError(f"no device dtype for type {ty}")

    def _const(self, op: Operation):
        ty, val = parse_constant(op.attrs["value"])
        s = self.sess
        if ty in ("HostFloat64Tensor", "Float64Tensor"):
            return s.from_host(FLOAT64, np.array(val, dtype=np.float64))
        if ty in _RING_TYPES:
            return s.from_ints(_RING_TYPES[ty], val)
        if ty in _BIT_TYPES:
            return s.from_bits(np.array(val, dtype=np.uint8))
        return val  # shapes, strings, keys, seeds, scalars

    def _max_value(self, op: Operation):
        return int(op.attrs["max_value"]) if "max_value" in op.attrs else None

    # -- one operation ----------------------------------------------------------------------------
    def _execute(self, op: Operation, x: List[object]):
        s, k = self.sess, op.kind
        if k == "PrfKeyGen":
            return s.gen_key()
        if k == "DeriveSeed":
            return s.derive_seed(attr_hex16(op, "sync_key"), x[0])
        if k == "Constant":
            return self._const(op)
        if k == "Input":
            raise ExecutionError("Input ops need `arguments`")
        if k in ("Identity",):
            return x[0]
        if k == "RingFixedpointEncode":
            if attr_int(op, "scaling_base") != 2:
                raise ExecutionError("RingFixedpointEncode: only scaling_base = 2")
            return s.fixedpoint_encode(self._dtype_of(op.output_type), attr_int(op, "scaling_exp"), x[0])
        if k == "RingFixedpointDecode":
            if attr_int(op, "scaling_base") != 2:
                raise ExecutionError("RingFixedpointDecode: only scaling_base = 2")
            return s.fixedpoint_decode(attr_int(op, "scaling_exp"), x[0])
        if k == "Shape":
            return tuple(x[0].shape)
        if k == "Fill":
            ty, val = parse_constant(op.attrs["value"])
            return s.fill(self._dtype_of(op.output_type), int(val), x[0])
        if k == "SampleSeeded":
            return s.sample_uniform_seeded(self._dtype_of(op.output_type), x[0], x[1], self._max_value(op))
        if k == "Sample":
            return s.sample_uniform(self._dtype_of(op.output_type), x[0], self._max_value(op))
        if k == "Add":
            return s.add(x[0], x[1])
        if k == "Sub":
            return s.sub(x[0], x[1])
        if k == "Mul":
            return s.mul(x[0], x[1])
        if k == "Dot":
            return s.dot(x[0], x[1])
        if k == "Neg":
            return s.neg(x[0])
        if k == "Shl":
            return s.shl(attr_int(op, "amount"), x[0])
        if k == "Shr":
            return s.shr(attr_int(op, "amount"), x[0])
        if k == "AddN":
            return s.add_n(x)
        if k == "Sum":
            axis = op.attrs.get("axis")
            return s.sum(x[0], None if axis in (None, "None") else int(axis))
        if k == "Xor":
            return s.xor(x[0], x[1])
        if k == "And":
            return s.and_(x[0], x[1])
        if k == "Or":
            return s.or_(x[0], x[1])
        if k == "BitDecompose":
            return s.bit_decompose(x[0]) if op.output_type in _BIT_TYPES else s.bit_decompose_ring(x[0])
        if k == "BitExtract":
            return s.bit_extract(attr_int(op, "bit_idx"), x[0])
        if k == "RingInject":
            return s.ring_inject(self._dtype_of(op.output_type), attr_int(op, "bit_idx"), x[0])
        if k == "ShlDim":
            return s.shl_dim(attr_int(op, "amount"), attr_int(op, "bit_length"), x[0])
        if k == "IndexAxis":
            return s.index_axis(attr_int(op, "axis"), attr_int(op, "index"), x[0])
        if k == "ExpandDims":
            return s.expand_dims(attr_int_list(op, "axis"), x[0])
        if k == "Squeeze":
            return s.squeeze(attr_int(op, "axis") if "axis" in op.attrs else None, x[0])
        if k == "Transpose":
            return s.transpose(x[0])
        if k == "Concat":
            return s.concat(attr_int(op, "axis"), x)
        if k == "Slice":
            info = attr_slice(op, "slice")
            if isinstance(x[0], tuple):  # SliceOp::shape_kernel (host/ops.rs:537-546)
                return tuple(x[0][info[0][0]:info[0][1]])
            return s.slice(info, x[0])
        if k == "Diag":
            return s.diag(x[0])
        if k == "Select":
            return s.select(attr_int(op, "axis"), x[1], x[0])
        if k == "Div":
            return s.div(x[0], x[1])
        if k == "Less":
            return s.less(x[0], x[1])
        if k == "Greater":
            return s.greater(x[0], x[1])
        if k == "Mux":
            return s.host_mux(x[0], x[1], x[2])
        if k == "Sign":
            return s.sign(x[0])
        if k == "Cast":
            return s.cast_ring128_to_ring64(x[0])
        if k == "Maximum":
            return s.host_maximum(x)
        if k == "RingFixedpointMean":
            axis = attr_int(op, "axis") if "axis" in op.attrs else None
            return s.fixedpoint_mean(axis, attr_int(op, "scaling_base"), attr_int(op, "scaling_exp"), x[0])
        if k == "RingFixedpointArgmax":
            return s.fixedpoint_argmax(attr_int(op, "axis"), x[0])
        if k == "Reshape":
            return s.reshape(x[0], x[1])
        if k == "Broadcast":
            return s.broadcast(x[1], x[0])
        if k == "Output":
            self.outputs[attr_str(op, "tag")] = x[0]
            return x[0]
        if k == "Save":
            self.saved[x[0]] = x[1]
            return None
        raise ExecutionError(f"operator {k} is not part of the host-dialect subset built here")

    # -- whole computation --------------------------------------------------------------------------
    def run(self, ops: List[Operation], arguments: Optional[Dict[str, object]] = None, keep_values: bool = True):
        arguments = arguments or {}
        for op in toposort(ops):
            if op.placement_kind != "Host":
                raise ExecutionError(f"{op.name}: only lowered graphs (@Host placements) can be replayed")
            mine = self.role is None or op.owners[0] == self.role
            if op.kind == "Send":
                if mine:
                    val = self.values[op.inputs[0]]
                    if self.comm is not None:
                        self.comm.send(val, attr_str(op, "receiver"))
                    else:
                        self.mailbox[op.attrs["rendezvous_key"]] = val
                    self.values[op.name] = None
                    self.executed += 1
                continue
            if op.kind == "Receive":
                if mine:
                    if self.comm is not None:
                        self.values[op.name] = self.comm.recv(attr_str(op, "sender"))
                    else:
                        self.values[op.name] = self.mailbox.pop(op.attrs["rendezvous_key"])
                    self.executed += 1
                continue
            if not mine:
                continue
            if op.name in self.overrides:
                self.values[op.name] = self.overrides[op.name]
            elif op.kind == "Input":
                self.values[op.name] = arguments[attr_str(op, "arg_name")]
            else:
                self.values[op.name] = self._execute(op, [self.values[i] for i in op.inputs])
            self.executed += 1
        return self.outputs
