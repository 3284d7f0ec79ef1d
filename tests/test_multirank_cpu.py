"""CPU tests (gloo, world_size 2) of the N>1 path of bench.py: one independent session per rank, no data-path
collective; the only cross-rank steps are the barrier, the max-over-ranks device time and the summed launch
count, then the whole-job aggregate computed on rank 0."""
import os
import sys

import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    import bench
    dist = bench.dist_setup(world, rank, backend="gloo")
    dist.barrier()
    ms, ms_e2e = bench.reduce_max(dist, [10.0 + 5 * rank, 30.0 - 7 * rank], rank)
    launches = bench.reduce_sum_int(dist, 100 + rank, rank)
    value = bench.aggregate_value(bench.ring_ops_per_step(8, 8, 8), steps=4, world=world, max_ms=ms)
    q.put((rank, ms, ms_e2e, launches, value))
    dist.barrier()
    dist.destroy_process_group()


def test_world2_reduction_and_aggregate():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29611 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ms, ms_e2e, launches, value in results:
        assert ms == 15.0 and ms_e2e == 30.0        # max over ranks
        assert launches == 201                      # summed
        ops = 3 * 2 * 2 * 8 * 8 * 8
        assert value == pytest.approx(world * ops * 4 / 15e-3)


def test_single_rank_helpers_are_identity():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.dist_setup(1, 0) is None
    assert bench.reduce_max(None, [1.5, 2.5]) == [1.5, 2.5]
    assert bench.reduce_sum_int(None, 7) == 7


def _party_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank), "WORLD_SIZE": str(world)})
    import torch.distributed as dist
    from moose_b200.executor import NcclComm
    dist.init_process_group("gloo")
    comm = NcclComm(dist, None, {"alice": 0, "bob": 1}, "cpu")
    seed, shape = bytes(range(16)), (65536, 512)
    if rank == 0:
        comm.send(seed, "bob")
        comm.send(shape, "bob")
        comm.send(None, "bob")
        got = comm.recv("bob")
    else:
        got = (comm.recv("alice"), comm.recv("alice"), comm.recv("alice"))
        comm.send(bytes(reversed(seed)), "alice")
    # the raw-buffer transport the native executor calls (mb_transport_fn): header + host payload, then a poll
    import ctypes
    hdr = (ctypes.c_int64 * 10)(*([1, 0, 0, 0, 0, 0, 0, 0, 0, 16] if rank == 0 else [0] * 10))
    payload = (ctypes.c_uint8 * 16)(*(range(100, 116) if rank == 0 else [0] * 16))
    peer = b"bob" if rank == 0 else b"alice"
    mode = 1 if rank == 0 else 0
    rc = comm.transport(None, mode, peer, ctypes.addressof(hdr), ctypes.sizeof(hdr), 0)
    rc |= comm.transport(None, mode, peer, ctypes.addressof(payload), 16, 0)
    done = ctypes.c_int64(-1)
    rc |= comm.transport(None, 3, b"", ctypes.addressof(done), 8, 0)
    got = (got, rc, list(hdr), bytes(payload), done.value)
    q.put((rank, got))
    dist.barrier()
    dist.destroy_process_group()


def test_party_wire_format_over_gloo():
    """Send/Receive of the non-tensor value kinds of a lowered graph (HostSeed / HostPrfKey, HostShape, HostUnit)
    between two parties: header + payload framing of moose_b200.executor.NcclComm."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29811 + os.getpid() % 100
    procs = [ctx.Process(target=_party_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[1][0] == (bytes(range(16)), (65536, 512), None)
    assert res[0][0] == bytes(reversed(range(16)))
    for r in (0, 1):  # transport callback: both sides succeed, the receiver sees header and payload, no device sends pending
        assert res[r][1] == 0 and res[r][2] == [1, 0, 0, 0, 0, 0, 0, 0, 0, 16] and res[r][3] == bytes(range(100, 116)) and res[r][4] == 0


def _shard_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    import numpy as np
    import bench
    dist = bench.dist_setup(world, rank, backend="gloo")
    batch, cols = 72, 3           # 9 blocks of 8 rows over 2 ranks: unequal slices (40 / 32 rows)
    lo, hi = bench.shard_rows(batch, world, rank, granule=8)
    full = np.arange(batch * cols, dtype=np.float64).reshape(batch, cols)
    rows_of = [b - a for a, b in (bench.shard_rows(batch, world, g, granule=8) for g in range(world))]
    got = bench.gather_rows(dist, full[lo:hi] * 2.0, rank, rows_of)        # each rank "computes" its rows of ONE batch
    ms = bench.reduce_max(dist, [20.0 if rank == 0 else 25.0], rank)[0]
    q.put((rank, (lo, hi), None if got is None else got.tolist(), batch * 4 / (ms * 1e-3)))
    dist.barrier()
    dist.destroy_process_group()


def test_world2_batch_sharding_and_output_gather():
    """one batch split by rows over the ranks (strong scaling), revealed outputs gathered on rank 0
    in row order, rows/s = whole batch x steps / max-over-ranks time."""
    import numpy as np
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29511 + os.getpid() % 90
    procs = [ctx.Process(target=_shard_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {r[0]: r[1:] for r in (q.get(timeout=120) for _ in range(world))}
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][0] == (0, 40) and res[1][0] == (40, 72)
    assert res[1][1] is None
    assert np.array_equal(np.array(res[0][1]), np.arange(72 * 3, dtype=np.float64).reshape(72, 3) * 2.0)
    assert res[0][2] == pytest.approx(72 * 4 / 25e-3)


def test_shard_rows_granule():
    sys.path.insert(0, ROOT)
    import bench
    rows = [bench.shard_rows(1000000, 8, g) for g in range(8)]
    assert rows[0] == (0, 125056) and rows[-1][1] == 1000000
    assert all(a[1] == b[0] for a, b in zip(rows, rows[1:]))                    # contiguous, in rank order
    assert all((hi - lo) % 64 == 0 and abs((hi - lo) - 125000) <= 64 for lo, hi in rows)
    assert [bench.shard_rows(100, 4, g, granule=64) for g in range(4)] == [(0, 64), (64, 100), (100, 100), (100, 100)]
