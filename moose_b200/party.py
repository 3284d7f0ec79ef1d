"""Party-per-GPU execution of a networked lowered computation (BASELINE config 5 mode): one process per
party, each on its own GPU, every Send/Receive pair of the graph (inserted by the reference's networking
pass, moose/src/compilation/networking.rs:75-118) carried by NCCL send/recv between the GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 3 --master-addr 127.0.0.1 --master-port P \
        -m moose_b200.party --graph tests/golden/dotprod_networked.json.gz [--repeat R]

Roles are mapped to ranks in order of first appearance in the graph.  Every party keeps its own session
(same session id, its own entropy), exactly like three comet workers of the reference.
"""
import argparse
import gzip
import json
import os
import time

import numpy as np


def roles_of(ops):
    roles = []
    for op in ops:
        for o in op.owners:
            if o not in roles:
                roles.append(o)
    return roles


def run_fused_mlp(args, group=None):
    """BASELINE config 5 on the party runtime: alice / bob / carole on three GPUs, reshares and openings over NCCL or
    the peer-memory rings.  Stand-alone (torchrun with 3 ranks: initialises and destroys the process group, prints the
    record) or, with `group` = a 3-rank subgroup of an initialised job (bench.py --gpus N >= 4 runs the parties on ranks
    0..2), returns the record on the group's first rank."""
    import torch
    import torch.distributed as dist
    import moose_b200 as mb
    from moose_b200.executor import NcclComm
    from moose_b200.party_rt import PartyMLP, PartySession

    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    own_group = group is None
    if own_group:
        if int(os.environ.get("WORLD_SIZE", "1")) != 3:
            raise SystemExit("the party runtime needs exactly 3 ranks")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    rank, world = dist.get_rank(group), 3
    src0 = dist.get_global_rank(group, 0) if group is not None else 0
    roles = ["player0", "player1", "player2"]
    sess = mb.Session(local_rank, args.session_id, None)
    native = None
    lanes = []          # --lanes K: K independent party sessions (own stream, own rings) working on B / K rows each
    if args.native_nccl:
        from moose_b200.party_rt import NativeNccl
        box = [NativeNccl.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=src0, group=group)          # bootstrap only: the unique id
        native = NativeNccl(sess, box[0], rank, roles)
        party = PartySession(sess, rank, native, roles)
    elif args.peer_memory:
        from moose_b200.party_rt import PeerMemory
        for lane in range(args.lanes):
            ls = sess if lane == 0 else mb.Session(local_rank, args.session_id, None)
            # the other lanes' spin waits stay resident on some SMs: leave those free so PRF grids run in one wave
            reserve = getattr(args, "reserve_sms", None)
            ls.reserve_sms(max(0, args.lanes - 1) if reserve is None else reserve)
            pm = PeerMemory(ls, rank, roles, args.ring_mib << 20)
            handles = [None] * world
            dist.all_gather_object(handles, pm.handle_bytes(), group=group)   # bootstrap only: the inbox handles
            pm.connect(handles)
            lanes.append((ls, pm, PartySession(ls, rank, pm, roles)))
        dist.barrier(group=group)
        native, party = lanes[0][1], lanes[0][2]
    else:
        comm = NcclComm(dist, sess, {r: i for i, r in enumerate(roles)}, local_rank)
        party = PartySession(sess, rank, comm.transport, roles)
    if not lanes:
        lanes = [(sess, native, party)]
    dims, B = args.dims, args.mlp
    if B % len(lanes):
        raise SystemExit("--lanes must divide the batch")
    rows = B // len(lanes)
    ws = [np.random.default_rng(4 + l).normal(0, dims[l] ** -0.5, (dims[l], dims[l + 1])) for l in range(len(dims) - 1)]
    bs = [np.random.default_rng(40 + l).normal(0, 0.1, dims[l + 1]) for l in range(len(dims) - 1)]
    mlps = [PartyMLP(pt, ws, bs, mb.RING64, 23) for _, _, pt in lanes]
    x = np.random.default_rng(3).uniform(0, 1, (B, dims[0]))
    xfs = [ls.fixedpoint_encode(mb.RING64, 23, ls.from_host(mb.FLOAT64, x[i * rows:(i + 1) * rows])) if rank == 0 else None
           for i, (ls, _, _) in enumerate(lanes)]

    def run_once():
        if len(lanes) == 1:
            outs = [mlps[0].forward(xfs[0], rows)]
        else:
            # one host thread per lane: a lane's host side waits for its peers' host messages (keys, seeds), which
            # arrive at the pace of the slowest party, so lanes enqueued from one thread run one after the other
            import threading
            outs, errs = [None] * len(lanes), []

            def work(i):
                try:
                    outs[i] = mlps[i].forward(xfs[i], rows)
                except BaseException as e:  # noqa: BLE001
                    errs.append(e)
            ts = [threading.Thread(target=work, args=(i,)) for i in range(len(lanes))]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
            if errs:
                raise errs[0]
        for ls, _, pt in lanes:
            pt.flush()
            ls.sync()
        return outs

    dist.barrier(group=group)
    outs = run_once()          # untimed: connections, scratch growth
    sent0 = sum(pt.bytes_sent for _, _, pt in lanes)
    st0 = party.stats()
    dist.barrier(group=group)
    t0 = time.perf_counter()
    for _ in range(args.repeat):
        outs = run_once()
    dt = (time.perf_counter() - t0) / args.repeat
    dist.barrier(group=group)
    trace = None
    if getattr(args, "trace", False):      # one more inference with an event behind every launch: busy vs waiting per party
        for ls, _, _ in lanes:
            ls.trace_enable(True)
        run_once()
        trace = [ls.trace_read() for ls, _, _ in lanes]
        for ls, _, _ in lanes:
            ls.trace_enable(False)
        dist.barrier(group=group)
    out = outs[0]
    sent = torch.tensor([sum(pt.bytes_sent for _, _, pt in lanes) - sent0], dtype=torch.int64, device=f"cuda:{local_rank}")
    dist.all_reduce(sent, group=group)
    err = None
    if out is not None:
        h = x[:8]
        for l in range(len(dims) - 1):
            h = h @ ws[l] + bs[l]
            if l < len(dims) - 2:
                h = np.maximum(h, 0)
        err = float(np.abs(out.to_host()[:8] - h).max())
    st1 = party.stats()
    mine = {k: (st1[k] - st0[k]) / args.repeat for k in st1}
    if trace is not None:
        mine["trace_ms"] = [{k: round(v["ms"], 3) for k, v in sorted(t.items(), key=lambda kv: -kv[1]["ms"])} for t in trace]
    gathered = [None] * world
    dist.all_gather_object(gathered, mine, group=group)
    record = None
    if rank == 0:
        record = ({"mode": "party-per-gpu (party runtime, fused party-sliced kernels)", "parties": roles,
                          "per_party": gathered,
                          "transport": ("peer-memory rings over NVLink, stream-ordered (csrc/peer_transport.cu)" if args.peer_memory
                                        else "ncclSend/ncclRecv from C++ (native)" if native
                                        else "nccl send/recv (torch.distributed)"),
                          "workload": "mlp_" + "-".join(map(str, dims)) + f"_relu_ring64_fixed(14,23)_batch{B}",
                          "lanes": len(lanes), "rows_per_s": B / dt, "seconds_per_run": dt, "payload_bytes_per_run": int(sent[0]) // args.repeat,
                          "max_abs_err_first_rows": err})
        if own_group:
            print(json.dumps(record), flush=True)
    for _, _, pt in lanes:
        pt.close()
    dist.barrier(group=group)          # nobody writes into an inbox that is about to be freed
    for ls, tr, _ in lanes:
        if tr is not None:
            tr.close()
        if ls is not sess:
            ls.close()
    sess.close()
    if own_group:
        dist.destroy_process_group()
    return record


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--graph", default=None, help=".moose text or the .json(.gz) re-encoding of tests/golden")
    ap.add_argument("--mlp", type=int, default=None, metavar="BATCH",
                    help="BASELINE config 5: emit the lowered + networked MLP graph (moose_b200/lowering.py) for --dims on "
                         "Ring64 fixed(14,23) and run it with one party per GPU on a synthetic batch")
    ap.add_argument("--dims", type=int, nargs="+", default=[784, 512, 512, 10])
    ap.add_argument("--native-nccl", action="store_true",
                    help="with --fused: the library's own NCCL transport (ncclSend / ncclRecv from C++) instead of "
                         "torch.distributed P2P")
    ap.add_argument("--peer-memory", action="store_true",
                    help="the peer-memory transport (rings in the receivers' HBM written over NVLink, every "
                         "transfer ordered on the party's stream, no host synchronisation per transfer)")
    ap.add_argument("--lanes", type=int, default=1,
                    help="with --peer-memory: split the batch over this many independent party sessions (own stream and "
                         "rings each); one lane's waits for its peers are filled by another lane's kernels")
    ap.add_argument("--ring-mib", type=int, default=1024, help="with --peer-memory: bytes of one ring, in MiB")
    ap.add_argument("--trace", action="store_true", help="with --fused: per-launcher device time of one inference per party")
    ap.add_argument("--reserve-sms", type=int, default=None,
                    help="with --lanes: SMs the PRF kernels leave free (default lanes - 1: one resident wait per other lane)")
    ap.add_argument("--fused", action="store_true",
                    help="with --mlp: run the party runtime (csrc/party.cpp: party-sliced fused kernels, msb-pruned adder, "
                         "seeds on the wire) instead of the per-op lowered graph")
    ap.add_argument("--session-id", default="b200-party")
    ap.add_argument("--repeat", type=int, default=1)
    ap.add_argument("--scale", type=int, nargs=3, metavar=("M", "K", "N"), default=None,
                    help="replace the two HostFloat64Tensor constants of the graph by U(-1,1) matrices [M,K] and [K,N]")
    args = ap.parse_args()
    if args.lanes > 1:
        # every lane's context owns two streams and parks spinning wait kernels on them; with more such streams than
        # hardware queues (8 by default) a signal can be queued behind another lane's wait.  Must be set before the
        # process creates its CUDA context.  (A lead for the 4-lane hang of DESIGN.md section 6, not re-measured.)
        os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

    import torch
    import torch.distributed as dist
    import moose_b200 as mb
    from moose_b200 import textual as tx
    from moose_b200.executor import Executor, NcclComm

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.mlp is not None and args.fused:
        return run_fused_mlp(args)
    if args.mlp is not None:
        from moose_b200.lowering import Lowering, networking_pass, prune_pass
        ops = networking_pass(prune_pass(Lowering(64).mlp_program(len(args.dims) - 1, 23)))
    elif args.graph is None:
        raise SystemExit("--graph or --mlp is required")
    else:
        raw = gzip.open(args.graph).read().decode() if args.graph.endswith(".gz") else open(args.graph).read()
        ops = tx.loads_json(raw) if args.graph.endswith((".json", ".json.gz")) else tx.parse_computation(raw)
    roles = roles_of(ops)
    if world != len(roles):
        raise SystemExit(f"graph has {len(roles)} parties {roles}, launched with {world} ranks")
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    role = roles[rank]
    sess = mb.Session(local_rank, args.session_id, None)  # OS entropy, like PrfKeyGen in the reference
    if args.peer_memory:
        from moose_b200.party_rt import PeerMemory
        comm = PeerMemory(sess, rank, roles, args.ring_mib << 20)
        handles = [None] * world
        dist.all_gather_object(handles, comm.handle_bytes())   # bootstrap only: the inbox handles
        comm.connect(handles)
    else:
        comm = NcclComm(dist, sess, {r: i for i, r in enumerate(roles)}, local_rank)
    # the two float constants, left operand first: decided by the constant's own shape ([1,3] . [3,1] in the exemplar)
    consts = [op for op in ops if op.kind == "Constant" and "Float64Tensor" in op.attrs.get("value", "")]
    consts.sort(key=lambda op: np.array(tx.parse_constant(op.attrs["value"])[1]).shape[0])
    overrides = {}
    mlp_ref = None
    if args.mlp is not None:
        # synthetic model and batch (BASELINE C4 shapes by default); every party binds its copy of the mirrored weights
        dims, rng = args.dims, np.random.default_rng(3)
        x = rng.uniform(0, 1, (args.mlp, dims[0]))
        h = x[:8]
        for l in range(len(dims) - 1):
            w = np.random.default_rng(4 + l).normal(0, dims[l] ** -0.5, (dims[l], dims[l + 1]))
            b = np.random.default_rng(40 + l).normal(0, 0.1, (1, dims[l + 1]))
            overrides[f"w{l}"], overrides[f"b{l}"] = sess.from_host(mb.FLOAT64, w), sess.from_host(mb.FLOAT64, b)
            h = h @ w + b
            if l < len(dims) - 2:
                h = np.maximum(h, 0)
        if rank == 0:
            overrides["x"] = sess.from_host(mb.FLOAT64, x)
        mlp_ref = h
    if args.scale:
        M, K, N = args.scale
        shapes = [(M, K), (K, N)]
        for i, op in enumerate(consts[:2]):
            if op.owners[0] == role:
                overrides[op.name] = sess.from_host(mb.FLOAT64, np.random.default_rng(100 + i).uniform(-1, 1, shapes[i]))
    dist.barrier()
    # one untimed run (NCCL connection setup, scratch growth), then the timed ones
    warm = Executor(sess, role=role, comm=comm, overrides=overrides, keep_values=False)
    warm.run(ops)
    warm.close()
    sess.sync()
    comm.bytes_sent = 0
    dist.barrier()
    outputs, t0 = {}, time.perf_counter()
    ex = None
    for _ in range(args.repeat):
        if ex is not None:
            ex.close()
        ex = Executor(sess, role=role, comm=comm, overrides=overrides, keep_values=False)
        outputs = ex.run(ops)
        sess.sync()
    dt = (time.perf_counter() - t0) / args.repeat
    dist.barrier()
    sent = torch.tensor([comm.bytes_sent], dtype=torch.int64, device=f"cuda:{local_rank}")
    dist.all_reduce(sent)
    result = {}
    for k, v in outputs.items():
        if isinstance(v, mb.Tensor):
            h = v.to_host()
            if mlp_ref is not None:
                result[k] = {"shape": list(h.shape), "max_abs_err_first_rows": float(np.abs(h[:8] - mlp_ref).max())}
            elif args.scale:  # check the first rows against a float matmul of the regenerated inputs
                M, K, N = args.scale
                a = np.random.default_rng(100).uniform(-1, 1, (M, K))[:4]
                b = np.random.default_rng(101).uniform(-1, 1, (K, N))
                result[k] = {"shape": list(h.shape), "max_abs_err_first_rows": float(np.abs(h[:4] - a @ b).max())}
            else:
                result[k] = h.tolist()
        else:
            result[k] = v
    gathered = [None] * world
    dist.all_gather_object(gathered, {"role": role, "executed": ex.executed, "transport_s": round(ex.transport_seconds, 4),
                                      "stream_wait_s": round(ex.sync_seconds, 4), "outputs": result})
    if rank == 0:
        extra = {} if args.mlp is None else {"workload": "mlp_" + "-".join(map(str, args.dims)) + f"_relu_ring64_fixed(14,23)_batch{args.mlp}",
                                              "rows_per_s": args.mlp / dt}
        print(json.dumps({"mode": "party-per-gpu", "parties": roles,
                          "transport": "peer-memory rings over NVLink (csrc/peer_transport.cu)" if args.peer_memory else "nccl send/recv",
                          **extra,
                          "ops": len(ops), "seconds_per_run": dt, "payload_bytes_per_run": int(sent[0]) // args.repeat,
                          "per_party": gathered}), flush=True)
    if ex is not None:
        ex.close()
    dist.barrier()          # nobody writes into an inbox that is about to be freed
    if args.peer_memory:
        comm.close()
    sess.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
