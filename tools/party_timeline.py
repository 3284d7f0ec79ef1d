"""Per-step device time of one party of the MLP (config 5) on the peer-memory transport: CUDA events on the party's
stream between the protocol steps (a step's time includes waiting for its peers).
torchrun --nproc-per-node 3 tools/party_timeline.py [BATCH]"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import moose_b200 as mb  # noqa: E402
from moose_b200.party_rt import PartyMLP, PartySession, PeerMemory  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
roles = ["player0", "player1", "player2"]
sess = mb.Session(rank, "timeline", None)
pm = PeerMemory(sess, rank, roles, 1 << 30)
hs = [None] * world
dist.all_gather_object(hs, pm.handle_bytes())
pm.connect(hs)
dist.barrier()
p = PartySession(sess, rank, pm, roles)
dims = [784, 512, 512, 10]
ws = [np.random.default_rng(4 + l).normal(0, dims[l] ** -0.5, (dims[l], dims[l + 1])) for l in range(3)]
bs = [np.random.default_rng(40 + l).normal(0, 0.1, dims[l + 1]) for l in range(3)]
mlp = PartyMLP(p, ws, bs, mb.RING64, 23)
x = np.random.default_rng(3).uniform(0, 1, (B, dims[0]))
xf = sess.fixedpoint_encode(mb.RING64, 23, sess.from_host(mb.FLOAT64, x)) if rank == 0 else None
stream = torch.cuda.ExternalStream(sess.stream, device=torch.device("cuda", rank))
marks = []


def mark(label):
    ev = torch.cuda.Event(enable_timing=True)
    ev.record(stream)
    marks.append((label, ev))


def forward():
    s = sess
    mark("start")
    h = p.share(0, xf if rank == 0 else None, mb.RING64, (B, dims[0]))
    mark("share_x")
    for i in range(3):
        w = p.share(0, mlp.w[i] if rank == 0 else None, mb.RING64, mlp.w[i].shape)
        mark(f"share_w{i}")
        d = p.dot(h, w)
        mark(f"dot{i}")
        t = p.trunc_pr(23, d)
        mark(f"trunc{i}")
        z = p.add_public(t, mlp.b[i])
        if i < 2:
            shape = (B, dims[i + 1])
            zeros = p.share(1, s.fill(mb.RING64, 0, shape) if rank == 1 else None, mb.RING64, shape)
            mark(f"share_zeros{i}")
            z = p.maximum([zeros, z])
            mark(f"maximum{i}")
        h = z
    out = p.reveal(0, h)
    mark("reveal")
    return out


for it in range(3):
    marks.clear()
    dist.barrier()
    forward()
    p.flush()
    sess.sync()
steps = {marks[i][0]: round(marks[i - 1][1].elapsed_time(marks[i][1]), 3) for i in range(1, len(marks))}
steps["total"] = round(marks[0][1].elapsed_time(marks[-1][1]), 3)
allsteps = [None] * world
dist.all_gather_object(allsteps, steps)
if rank == 0:
    for k in steps:
        print(f"{k:14s}" + "".join(f"{a[k]:10.3f}" for a in allsteps), flush=True)
dist.barrier()
p.close()
dist.barrier()
pm.close()
sess.close()
dist.destroy_process_group()
