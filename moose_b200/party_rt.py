"""One party of the replicated protocols per process / GPU: Python face of csrc/party.cpp (mb_party_* entry points).

Each of the three parties creates a `PartySession` on its own GPU with the same session id and issues the same
sequence of calls (SPMD); reshares, the dealer's messages and openings travel through the transport -- NCCL
send/recv between the parties' GPUs (`NcclComm.transport` of moose_b200/executor.py), or the in-process
`LocalTransport` below for tests that run the three parties as threads on one GPU.
"""
from __future__ import annotations

import ctypes
import queue
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import FLOAT64, RING64, RING128, Session, Tensor, _shape_arr, lib
from .executor import TRANSPORT_FN, SendRecv


class NativeNccl:
    """The library's own NCCL transport (csrc/nccl_transport.cpp).  `unique_id` (128 bytes) comes from
    NativeNccl.unique_id() on one rank and reaches the others through the host's bootstrap (torch.distributed here)."""

    def __init__(self, sess: Session, unique_id: bytes, rank: int, role_names: Sequence[str]):
        names = (ctypes.c_char_p * len(role_names))(*[n.encode() for n in role_names])
        self._names, self.sess = names, sess
        self.handle = ctypes.c_void_p()
        sess._check(lib().mb_nccl_transport_create(sess.ctx, unique_id, rank, len(role_names), names, ctypes.byref(self.handle)))
        self.fn = lib().mb_nccl_transport_fn()

    @staticmethod
    def unique_id() -> bytes:
        buf = (ctypes.c_uint8 * 128)()
        if lib().mb_nccl_unique_id(buf) != 0:
            raise RuntimeError("mb_nccl_unique_id failed: " + (lib().mb_last_error(None) or b"").decode())
        return bytes(buf)

    def close(self):
        if self.handle:
            lib().mb_nccl_transport_free(self.handle)
            self.handle = ctypes.c_void_p()


class PeerMemory:
    """The library's peer-memory transport (csrc/peer_transport.cu): rings in the receivers' HBM written over NVLink,
    every transfer enqueued on the context's stream.  Start-up in two steps: create on every rank, hand the
    128-byte `handle_bytes()` of all ranks (rank order) to `connect`.  Call `close()` only after a barrier."""

    def __init__(self, sess: Session, rank: int, role_names: Sequence[str], ring_bytes: int = 0):
        names = (ctypes.c_char_p * len(role_names))(*[n.encode() for n in role_names])
        self._names, self.sess = names, sess
        self.handle = ctypes.c_void_p()
        sess._check(lib().mb_peer_transport_create(sess.ctx, rank, len(role_names), names, ring_bytes, ctypes.byref(self.handle)))
        self.fn = lib().mb_peer_transport_fn()

    def handle_bytes(self) -> bytes:
        buf = (ctypes.c_uint8 * 128)()
        self.sess._check(lib().mb_peer_transport_handle(self.handle, buf))
        return bytes(buf)

    def connect(self, handles: Sequence[bytes]):
        blob = b"".join(handles)
        self.sess._check(lib().mb_peer_transport_connect(self.handle, blob))

    def close(self):
        if self.handle:
            lib().mb_peer_transport_free(self.handle)
            self.handle = ctypes.c_void_p()


class PShare:
    """This party's two shares (x_me, x_{me+1}) of a replicated tensor."""

    def __init__(self, party: "PartySession", handle: int):
        self.party, self.h = party, handle

    def __del__(self):
        try:
            if self.h and self.party._p:
                lib().mb_pshare_free(self.party._p, self.h)
        except Exception:
            pass
        self.h = None

    def share(self, i: int) -> Tensor:
        return Tensor(self.party.sess, lib().mb_pshare_tensor(self.h, i), owned=False)


class PartySession:
    def __init__(self, sess: Session, me: int, transport, role_names: Sequence[str] = ("player0", "player1", "player2")):
        """`transport`: a callable with the mb_transport_fn contract (mode, peer, buf, nbytes, on_device)."""
        self.sess, self.me = sess, me
        names = (ctypes.c_char_p * 3)(*[n.encode() for n in role_names])
        self._names = names
        self._p = ctypes.c_void_p()
        if isinstance(transport, (NativeNccl, PeerMemory)):   # C function + handle: no interpreter in the transfer loop
            self._cb, user = ctypes.c_void_p(transport.fn), transport.handle
        else:
            self._cb, user = TRANSPORT_FN(transport), None
        self._transport = transport
        sess._check(lib().mb_party_create(sess.ctx, me, self._cb, user, names, ctypes.byref(self._p)))

    def close(self):
        if self._p:
            lib().mb_party_free(self._p)
            self._p = ctypes.c_void_p()

    @property
    def bytes_sent(self) -> int:
        return int(lib().mb_party_bytes_sent(self._p))

    def stats(self) -> Dict[str, float]:
        f = lib().mb_party_stat
        return {"transport_s": f(self._p, 0), "stream_wait_s": f(self._p, 1), "tensor_sends": int(f(self._p, 2))}

    def flush(self):
        self.sess._check(lib().mb_party_flush(self._p))

    def _s(self, rc, out) -> PShare:
        self.sess._check(rc)
        return PShare(self, out.value)

    def share(self, owner: int, x: Optional[Tensor], dtype: int, shape) -> PShare:
        arr, rank = _shape_arr(shape)
        out = ctypes.c_void_p()
        return self._s(lib().mb_party_share(self._p, owner, x.h if x is not None else None, dtype, rank, arr, ctypes.byref(out)), out)

    def reveal(self, receiver: int, x: PShare) -> Optional[Tensor]:
        out = ctypes.c_void_p()
        self.sess._check(lib().mb_party_reveal(self._p, receiver, x.h, ctypes.byref(out)))
        return Tensor(self.sess, out.value) if out.value else None

    def _bin(self, fn, x, y) -> PShare:
        out = ctypes.c_void_p()
        return self._s(fn(self._p, x.h, y.h, ctypes.byref(out)), out)

    def add(self, x, y): return self._bin(lib().mb_party_add, x, y)
    def sub(self, x, y): return self._bin(lib().mb_party_sub, x, y)
    def mul(self, x, y): return self._bin(lib().mb_party_mul, x, y)
    def dot(self, x, y): return self._bin(lib().mb_party_dot, x, y)
    def less(self, x, y): return self._bin(lib().mb_party_less, x, y)
    def add_public(self, x: PShare, pub: Tensor): return self._bin(lib().mb_party_add_public, x, pub)

    def trunc_pr(self, amount: int, x: PShare) -> PShare:
        out = ctypes.c_void_p()
        return self._s(lib().mb_party_trunc_pr(self._p, amount, x.h, ctypes.byref(out)), out)

    def msb(self, x: PShare) -> PShare:
        out = ctypes.c_void_p()
        return self._s(lib().mb_party_msb(self._p, x.h, ctypes.byref(out)), out)

    def ring_inject(self, ring: int, bit_idx: int, x: PShare) -> PShare:
        out = ctypes.c_void_p()
        return self._s(lib().mb_party_ring_inject(self._p, ring, bit_idx, x.h, ctypes.byref(out)), out)

    def maximum(self, xs: Sequence[PShare]) -> PShare:
        arr = (ctypes.c_void_p * max(len(xs), 1))(*[x.h for x in xs])
        out = ctypes.c_void_p()
        return self._s(lib().mb_party_maximum(self._p, arr, len(xs), ctypes.byref(out)), out)


class PartyMLP:
    """pymoose MLPRegressor with ReLU = maximum([zeros, z]) (multilayer_perceptron_predictor.py:86-123), one party's
    side: x comes from party 0, the mirrored weights are shared from party 0's copy, biases are public, the ReLU zeros
    are shared by party 1, the logits are revealed to `reveal_to`."""

    def __init__(self, party: PartySession, weights: Sequence[np.ndarray], biases: Sequence[np.ndarray], ring: int = RING64,
                 frac: int = 23):
        self.p, self.ring, self.frac = party, ring, frac
        s = party.sess
        enc = lambda a: s.fixedpoint_encode(ring, frac, s.from_host(FLOAT64, np.asarray(a, dtype=np.float64)))
        self.w = [enc(w) for w in weights]
        self.b = [enc(np.asarray(b).reshape(1, -1)) for b in biases]

    def forward(self, x_fixed: Optional[Tensor], batch: int, reveal_to: int = 0) -> Optional[Tensor]:
        p, s = self.p, self.p.sess
        x = p.share(0, x_fixed if p.me == 0 else None, self.ring, (batch, self.w[0].shape[0]))
        n = len(self.w)
        for i in range(n):
            w = p.share(0, self.w[i] if p.me == 0 else None, self.ring, self.w[i].shape)
            z = p.add_public(p.trunc_pr(self.frac, p.dot(x, w)), self.b[i])
            if i < n - 1:
                shape = (batch, self.w[i].shape[1])
                zeros = p.share(1, s.fill(self.ring, 0, shape) if p.me == 1 else None, self.ring, shape)
                z = p.maximum([zeros, z])
            x = z
        out = p.reveal(reveal_to, x)
        return s.fixedpoint_decode(self.frac, out) if out is not None else None


class LocalTransport:
    """In-process transport for three parties running as threads on one GPU (tests): values are staged in a device
    tensor of the sender's session and handed over through a queue per directed pair."""

    def __init__(self, names: Sequence[str] = ("player0", "player1", "player2")):
        self.names = list(names)
        self.q: Dict[tuple, "queue.Queue"] = {(a, b): queue.Queue() for a in names for b in names if a != b}

    def endpoint(self, me: str, sess: Session):
        def fn(_user, mode, peer, buf, nbytes, on_device):
            try:
                if mode == 5:  # capabilities: none
                    ctypes.cast(buf, ctypes.POINTER(ctypes.c_int64))[0] = 0
                    return 0
                if mode in (2, 3):  # every send below is complete when it returns
                    ctypes.cast(buf, ctypes.POINTER(ctypes.c_int64))[0] = fn.sent
                    return 0
                if mode == 4:
                    d = ctypes.cast(buf, ctypes.POINTER(SendRecv))[0]
                    rc = fn(None, 1, d.send_peer, d.send_buf, d.send_bytes, 1)
                    return rc or fn(None, 0, d.recv_peer, d.recv_buf, d.recv_bytes, 1)
                other = peer.decode()
                if mode == 1:
                    if on_device:
                        # staged in whole 8-byte words: tensor allocations are padded to 16 bytes with zeros
                        stage = sess.alloc(RING64, ((nbytes + 7) // 8,))
                        sess._check(lib().mb_tensor_copy_from_device(sess.ctx, stage.h, buf))
                        sess.sync()
                        fn.sent += 1
                        self.q[(me, other)].put(stage)
                    else:
                        self.q[(me, other)].put(bytes((ctypes.c_uint8 * nbytes).from_address(buf)))
                else:
                    item = self.q[(other, me)].get(timeout=60)
                    if on_device:
                        # the staging tensor belongs to the sender's context: copy on ours after its copy is done
                        sess._check(lib().mb_tensor_copy_to_device(sess.ctx, item.h, buf))
                        sess.sync()
                    else:
                        ctypes.memmove(buf, item, nbytes)
                return 0
            except Exception as e:  # must not propagate through the C frames
                import sys
                print(f"LocalTransport error: {e!r}", file=sys.stderr, flush=True)
                return 1
        fn.sent = 0
        return fn
