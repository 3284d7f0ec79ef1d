import os, time, torch, torch.distributed as dist
import moose_b200 as mb
from moose_b200.executor import _DeviceBytes
rank = int(os.environ["RANK"]); torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
sess = mb.Session(rank, "p", None)
n = 268435456
def bench(label, make):
    bufs = [make() for _ in range(4)]
    for it in range(2):
        torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
        for b in bufs:
            (dist.send if rank == 0 else dist.recv)(b, 1 - rank)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    if rank == 0: print(label, f"{4*n/dt/1e9:.1f} GB/s", flush=True)
bench("torch caching allocator", lambda: torch.empty(n, dtype=torch.uint8, device="cuda"))
keep = []
def mk():
    t = sess.alloc(mb.RING128, (4096, 4096)); keep.append(t); sess.sync()
    return torch.as_tensor(_DeviceBytes(lib_ptr(t), n), device=torch.device("cuda", rank))
def lib_ptr(t): return mb.lib().mb_tensor_device_ptr(t.h)
bench("mempool (cudaMallocAsync)", mk)
dist.destroy_process_group()
