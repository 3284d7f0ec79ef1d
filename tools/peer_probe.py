"""Bandwidth of the peer-memory transport between three ranks: the ring exchange (mode 4: every rank sends to the
previous one and receives from the next) and one directed transfer, 256 MiB payloads.
torchrun --nproc-per-node 3 tools/peer_probe.py"""
import ctypes
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import moose_b200 as mb  # noqa: E402
from moose_b200.executor import TRANSPORT_FN, SendRecv
from moose_b200.party_rt import PeerMemory

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
names = [f"player{i}" for i in range(world)]
sess = mb.Session(rank, "probe", None)
pm = PeerMemory(sess, rank, names, int(os.environ.get("RING_MIB", "1024")) << 20)
hs = [None] * world
dist.all_gather_object(hs, pm.handle_bytes())
pm.connect(hs)
dist.barrier()
fn = TRANSPORT_FN(pm.fn)
N = 8
src = [sess.alloc(mb.RING128, (4096, 4096)) for _ in range(N)]
dst = [sess.alloc(mb.RING128, (4096, 4096)) for _ in range(N)]
nbytes = 4096 * 4096 * 16
ptr = lambda t: mb.lib().mb_tensor_device_ptr(t.h)
out = {}
for label in ("ring_exchange", "directed_0_to_1"):
    for it in range(3):
        sess.sync(); dist.barrier(); t0 = time.perf_counter()
        for i in range(N):
            if label == "ring_exchange":
                d = SendRecv(names[(rank - 1) % world].encode(), ptr(src[i]), nbytes, names[(rank + 1) % world].encode(), ptr(dst[i]), nbytes)
                assert fn(pm.handle, 4, b"", ctypes.byref(d), ctypes.sizeof(d), 1) == 0
            elif rank == 0:
                assert fn(pm.handle, 1, names[1].encode(), ptr(src[i]), nbytes, 1) == 0
            elif rank == 1:
                assert fn(pm.handle, 0, names[0].encode(), ptr(dst[i]), nbytes, 1) == 0
        t_enq = time.perf_counter() - t0
        sess.sync(); dt = time.perf_counter() - t0
    out[label] = {"GB_per_s_per_rank": N * nbytes / dt / 1e9, "enqueue_ms": t_enq * 1e3, "total_ms": dt * 1e3}
dist.barrier()
if rank == 0:
    print(json.dumps(out), flush=True)
dist.barrier()
pm.close()
sess.close()
dist.destroy_process_group()
