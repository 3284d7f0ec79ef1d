#!/usr/bin/env python
"""bench.py -- benchmark of the replicated-secret-sharing hot path (BASELINE.json: "MPC inference rows/sec
(LogReg/MLP) at 1/2/4/8 B200; Ring128 matmul GOP/s").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload all|matmul|mlp|logreg|prf]

One run measures three workloads and prints ONE JSON line (stdout); the line's top-level fields are the matmul
(BASELINE configs[1], the single-GPU config the metric is quoted on) and `workloads` carries the two predictors
(each record is also printed on stderr as it finishes):

  matmul  configs[1]  Replicated Ring128 fixed(24,40) 4096^3 dot + TruncPr(40), three parties on one B200.  A step =
          one DotOp::rep_rep_kernel + one TruncPrOp::rep_kernel.  Unit: ring GOP/s = 3 parties x 2 GEMMs x 2MNK / t
          (SURVEY.md 8d).  With --gpus N every rank runs its own copy (a 1-GPU config: "weak", replicas).
  mlp     configs[3]  MLP 784-512-512-10, ReLU = maximum([0, z]), Ring64 fixed(14,23), batch 65536, logits revealed,
          in the reference's op sequence (mirrored biases shared by player 0, fixedpoint/ops.rs:30-52).  Unit: rows/s.
  logreg  configs[2]  LinearClassifier, 512 features, 2 classes, sigmoid, Ring128 fixed(24,40), batch 10^6.  rows/s.

With --gpus N the two predictors split ONE batch over the N ranks (rank g runs all three parties on rows
[g*B/N, (g+1)*B/N), an independent session per GPU, no data-path collective): "strong" scaling.
`value` is device time of the step with inputs resident in HBM; `e2e` starts from a host float64 batch in pinned
memory and ends with the decoded outputs of ALL ranks in rank 0's host memory (H2D, encode, share, forward, reveal,
decode, NCCL gather of the outputs, D2H inside the timed region).

Each record carries `roofline` for its dominant kernel (from the library's per-kernel trace, a second pass of the same
steps with one CUDA event behind every launch), a `kernels` table of the launchers behind 95 % of the step, and at N = 1
a `cpu_baseline`: the oracle (a port of the reference's CPU kernels; the Rust reference cannot be built here) on a
bounded sample.  --impl reference times that port on the box's host cores for the same three workloads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAC = 40
SESSION_ID = "b200-bench"
ENTROPY_KEY = bytes(range(16))
MLP_DIMS = (784, 512, 512, 10)
MLP_FRAC, MLP_INT = 23, 14
LOGREG_FEATS, LOGREG_FRAC, LOGREG_INT = 512, 40, 24
L2_NOTE = "inputs larger than L2 (the shares of one step exceed 126 MB)"


def ring_ops_per_step(M, K, N):
    # 3 parties x 2 GEMMs x (M*N*K multiplies + M*N*K adds)
    return 3 * 2 * 2 * M * N * K


def fixed128(vals):
    """fixed(24,40) encoding as uint64[..., 2] = [lo, hi] (truncation toward zero, two's complement)."""
    scaled = np.trunc(vals * float(2 ** FRAC)).astype(np.int64)
    lo = scaled.astype(np.uint64)
    hi = np.where(scaled < 0, np.uint64(0xFFFFFFFFFFFFFFFF), np.uint64(0)).astype(np.uint64)
    return np.stack([lo, hi], axis=-1)


def log(msg):
    print(msg, file=sys.stderr, flush=True)


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU during the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None
        self.first = 0

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def mark(self):
        """A timed region starts here: rows read before count only if none arrives inside it (nvidia-smi needs
        100 ms and more to start, a short run can be over by then; the warm-up steps run the same load)."""
        self.first = len(self.rows)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def snapshot(self):
        """Clock summary of the rows since mark() (the sampler keeps running)."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        timed = self.rows[self.first:]
        source = "timed region"
        if not timed:
            timed, source = list(self.rows), "warm-up steps (same load; the timed region was shorter than one sample)"
        for r in timed:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max((float(r[3]) for r in timed if len(r) > 3 and r[3].replace('.', '', 1).isdigit()),
                                   default=None),
                "samples": len(sm), "sampled_during": source, "reasons": sorted(reasons)}

    def stop(self):
        out = self.snapshot()
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        return out


# ---- multi-rank helpers (one process per GPU under torchrun; CPU tests pass backend="gloo") --------------------------
def dist_setup(world, local_rank, backend=None):
    """One process per GPU (torchrun): returns torch.distributed or None.  `backend` defaults to nccl; the CPU
    tests of the multi-rank logic pass "gloo"."""
    if world <= 1:
        return None
    import torch
    import torch.distributed as dist
    backend = backend or "nccl"
    if backend == "nccl":
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    else:
        dist.init_process_group(backend)
    return dist


def _reduce_device(dist, local_rank):
    return f"cuda:{local_rank}" if dist.get_backend() == "nccl" else "cpu"


def reduce_max(dist, values, local_rank=0):
    """Max over ranks of each entry (device times of the timed regions)."""
    if dist is None:
        return [float(v) for v in values]
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=_reduce_device(dist, local_rank))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t]


def reduce_sum_int(dist, value, local_rank=0):
    if dist is None:
        return int(value)
    import torch
    t = torch.tensor([int(value)], dtype=torch.int64, device=_reduce_device(dist, local_rank))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return int(t[0])


def aggregate_value(units_per_rank_step, steps, world, max_ms):
    """Whole-job throughput: every rank processes the same number of units, time = max over ranks."""
    return world * units_per_rank_step * steps / (max_ms * 1e-3)


ROW_GRANULE = 64   # rows: keeps every rank's [rows, 2] and [rows, 512] tensors a multiple of 128 elements, the vector
                   # width of the fused bit-plane kernels (other row counts take the slower op-by-op composition)


def shard_rows(batch, world, rank, granule=ROW_GRANULE):
    """Rows [lo, hi) of ONE batch that rank `rank` of `world` processes: the batch is cut into blocks of `granule` rows
    and the blocks are dealt as evenly as possible (the first ranks get one more), so slices differ by at most one
    block (10^6 rows over 8 ranks: 125 056 / 124 992)."""
    blocks = -(-batch // granule)
    per, extra = divmod(blocks, world)
    first = rank * per + min(rank, extra)
    count = per + (1 if rank < extra else 0)
    return min(first * granule, batch), min((first + count) * granule, batch)


def gather_rows(dist, local, local_rank=0, rows_of=None):
    """Host-side gather of the revealed, decoded outputs: rank 0 gets the [B, cols] array in rank order, the others
    None.  `local` is this rank's [rows, cols] float64 array (host); `rows_of[g]` = rows of rank g when the slices are
    unequal (each rank then sends max(rows_of) rows, zero padded)."""
    if dist is None:
        return local
    import torch
    dev = _reduce_device(dist, local_rank)
    world = dist.get_world_size()
    rows_of = list(rows_of) if rows_of is not None else [local.shape[0]] * world
    pad = np.zeros((max(rows_of),) + local.shape[1:], dtype=local.dtype)
    pad[:local.shape[0]] = local
    t = torch.from_numpy(pad).to(dev)
    bufs = [torch.empty_like(t) for _ in range(world)] if dist.get_rank() == 0 else None
    dist.gather(t, bufs, dst=0)
    if bufs is None:
        return None
    return torch.cat([b[:r] for b, r in zip(bufs, rows_of)], dim=0).cpu().numpy()


class _DevView:
    """A library tensor's device memory as a __cuda_array_interface__ object (so torch / NCCL read it in place)."""

    def __init__(self, ptr, shape, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class OutputGather:
    """Per step: the decoded outputs of all ranks land in rank 0's pinned host memory.  N = 1: one D2H copy on the
    drainer context.  N > 1: an NCCL gather straight out of device memory, enqueued behind the drainer context's stream
    (torch ExternalStream), then rank 0's D2H copies of the gathered rows.  Slices may differ by one row granule, so every
    rank gathers max-rows rows (its tensor copied into a padded send buffer)."""

    def __init__(self, dist, drainer, local_rank, rows_of, cols, world, rank):
        self.dist, self.drainer, self.rank, self.world = dist, drainer, rank, world
        self.rows_of, self.rows, self.cols = list(rows_of), rows_of[rank], cols
        self.max_rows = max(rows_of)
        if dist is None:
            self.host = [drainer.pinned_empty((self.rows, cols), np.float64) for _ in range(2)]
            return
        import torch
        self.torch = torch
        self.dev = torch.device("cuda", local_rank)
        self.ext = torch.cuda.ExternalStream(drainer.stream, device=self.dev)
        self.send = torch.zeros((self.max_rows, cols), dtype=torch.float64, device=self.dev)
        if rank == 0:
            self.bufs = [torch.empty((self.max_rows, cols), dtype=torch.float64, device=self.dev) for _ in range(world)]
            self.host_t = [torch.empty((sum(rows_of), cols), dtype=torch.float64).pin_memory() for _ in range(2)]
            self.host = [t.numpy() for t in self.host_t]
        torch.cuda.synchronize(self.dev)

    def step(self, i, r):
        """r: this rank's decoded float64 tensor, already handed to the drainer context (r.sess = drainer)."""
        if self.dist is None:
            self.drainer.to_host_async(r, self.host[i % 2])
            return
        torch = self.torch
        with torch.cuda.stream(self.ext):
            t = torch.as_tensor(_DevView(r.device_ptr, (self.rows, self.cols)), device=self.dev)
            self.send[:self.rows].copy_(t)
            self.dist.gather(self.send, self.bufs if self.rank == 0 else None, dst=0)
            if self.rank == 0:
                lo = 0
                for b, n in zip(self.bufs, self.rows_of):
                    self.host_t[i % 2][lo:lo + n].copy_(b[:n], non_blocking=True)
                    lo += n

    def result(self, i):
        return self.host[i % 2] if (self.dist is None or self.rank == 0) else None

    def close(self):
        """Releases the torch buffers while the drainer's stream still exists (the caching allocator records an event
        on every stream a block was used on when the block is freed)."""
        if self.dist is None:
            return
        self.torch.cuda.synchronize(self.dev)
        if self.rank == 0:
            self.host = [h.copy() for h in self.host]
            self.bufs = self.host_t = None
        self.send = self.ext = None
        self.torch.cuda.synchronize(self.dev)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


def host_threads():
    # all host threads the process may use (torchrun exports OMP_NUM_THREADS=1, so ask the OS, not OpenMP)
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ---- synthetic inputs ---------------------------------------------------------------------------------
def mlp_model():
    ws = [np.random.default_rng(4 + i).normal(0, MLP_DIMS[i] ** -0.5, (MLP_DIMS[i], MLP_DIMS[i + 1])) for i in range(3)]
    bs = [np.random.default_rng(40 + i).normal(0, 0.1, MLP_DIMS[i + 1]) for i in range(3)]
    return ws, bs


def mlp_float(x, ws, bs):
    h = x
    for i in range(3):
        h = h @ ws[i] + bs[i]
        if i < 2:
            h = np.maximum(h, 0)
    return h


def logreg_model():
    coeffs = np.random.default_rng(2).normal(0, LOGREG_FEATS ** -0.5, (2, LOGREG_FEATS))
    intercepts = np.random.default_rng(22).normal(0, 0.1, (1, 2))
    return coeffs, intercepts


def logreg_float(x, coeffs, intercepts):
    return 1.0 / (1.0 + np.exp(-(x @ coeffs.T + intercepts)))


def fill_uniform(buf, seed, lo, hi):
    """Fills the (pinned) float64 matrix `buf` with U(lo, hi) in chunks (the LogReg batch is 4 GB)."""
    rng = np.random.default_rng(seed)
    rows = buf.shape[0]
    step = max(1, (1 << 24) // max(buf.shape[1], 1))
    for a in range(0, rows, step):
        b = min(rows, a + step)
        rng.random(out=buf[a:b])
        if (lo, hi) != (0.0, 1.0):
            buf[a:b] *= (hi - lo)
            buf[a:b] += lo


# ---- CPU legs (the oracle: a port of the reference's CPU kernels) ----------------------------------------------------
def cpu_matmul(size, threads, with_trunc=True):
    from oracle import moose_oracle as mo
    rng = np.random.default_rng(0)
    a = mo.Ring(128, fixed128(rng.uniform(-1, 1, (size, size))))
    b = mo.Ring(128, fixed128(rng.uniform(-1, 1, (size, size))))
    sess = mo.Session(SESSION_ID, ENTROPY_KEY)
    x, y = mo.rep_share(sess, a, 0), mo.rep_share(sess, b, 1)
    mo.DOT_THREADS = threads
    t0 = time.perf_counter()
    z = mo.rep_dot(sess, x, y)
    if with_trunc:
        z = mo.rep_trunc_pr(sess, FRAC, z)
    dt = time.perf_counter() - t0
    mo.DOT_THREADS = 1
    return ring_ops_per_step(size, size, size) / dt / 1e9, dt


cpu_baseline_run = cpu_matmul  # name used by round-1 callers


def cpu_mlp(batch, threads):
    from oracle import moose_oracle as mo, predictors_oracle as po
    ws, bs = mlp_model()
    x = np.random.default_rng(3).uniform(0, 1, (batch, MLP_DIMS[0]))
    sess = mo.Session(SESSION_ID + "-mlp", ENTROPY_KEY)
    mo.DOT_THREADS = threads
    t0 = time.perf_counter()
    _, dec = po.mlp(sess, x, ws, bs, bits=64, frac=MLP_FRAC, bias_via="share")
    dt = time.perf_counter() - t0
    mo.DOT_THREADS = 1
    assert np.allclose(dec, mlp_float(x, ws, bs), atol=5e-3)
    return batch / dt, dt


def cpu_logreg(batch, threads):
    from oracle import moose_oracle as mo, predictors_oracle as po
    coeffs, intercepts = logreg_model()
    x = np.random.default_rng(1).uniform(-1, 1, (batch, LOGREG_FEATS))
    sess = mo.Session(SESSION_ID + "-logreg", ENTROPY_KEY)
    mo.DOT_THREADS = threads
    t0 = time.perf_counter()
    _, dec = po.linear(sess, x, coeffs, intercepts, bits=128, frac=LOGREG_FRAC, integral=LOGREG_INT, post_transform="sigmoid")
    dt = time.perf_counter() - t0
    mo.DOT_THREADS = 1
    assert np.allclose(dec, logreg_float(x, coeffs, intercepts), atol=2e-2)
    return batch / dt, dt


PORT_NOTE = ("CPU = the oracle, a port of the reference's host kernels and protocol op sequence (the Rust reference cannot be "
             "built here); every host op single-threaded like the reference's ndarray kernels")


def workload_names(batch_mlp, batch_logreg, size):
    return {
        "matmul": f"replicated_ring128_matmul_{size}x{size}x{size}_truncpr40",
        "mlp": f"mlp_784-512-512-10_relu_ring64_fixed(14,23)_batch{batch_mlp}",
        "logreg": f"logreg_512features_2classes_sigmoid_ring128_fixed(24,40)_batch{batch_logreg}",
    }


def matmul_config(name):
    return {"workload": name, "fixed": "fixed(24,40)", "parties": "3 parties resident on each GPU; one independent session per GPU",
            "l2": L2_NOTE}


def predictor_config(name, world):
    return {"workload": name, "parties": "3 parties resident on each GPU; one independent session per GPU",
            "sharding": f"one batch split over {world} rank(s) by rows ({ROW_GRANULE}-row granule), no data-path collective; outputs gathered on rank 0",
            "op_sequence": "reference's logical graph (mirrored weights / biases shared by player 0)", "l2": L2_NOTE}


def run_reference(args, rank, world):
    """--impl reference: the CPU port on the box's host cores, same three workloads, each step a bounded sample."""
    if rank != 0:
        return
    threads = host_threads()
    names = workload_names(args.batch_mlp, args.batch_logreg, args.size)
    sub_steps, sub_warm = max(1, min(args.steps, 3)), min(args.warmup, 1)

    def timed(fn, arg, steps, warm, warm_arg):
        for _ in range(warm):
            fn(warm_arg, threads)
        times = []
        for _ in range(steps):
            _, dt = fn(arg, threads)
            times.append(dt)
        return times

    records = {}
    if args.workload in ("all", "mlp"):
        b = args.ref_batch_mlp
        times = timed(cpu_mlp, b, sub_steps, sub_warm, 32)
        v = b * len(times) / sum(times)
        sample = f"batch {b} of the {args.batch_mlp}-row workload per step, full MLP forward incl. sharing; dot over {threads} threads (OpenMP), other host ops single-threaded"
        records["mlp"] = {
            "impl": "reference", "metric": "mlp_inference_rows_per_s", "value": v, "unit": "rows/s", "steps": len(times),
            "warmup": sub_warm, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "dtype": "u64",
            "config": predictor_config(names["mlp"], world),
            "cpu_baseline": {"value": v, "unit": "rows/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        log(json.dumps(records["mlp"]))
    if args.workload in ("all", "logreg"):
        b = args.ref_batch_logreg
        times = timed(cpu_logreg, b, sub_steps, sub_warm, 64)
        v = b * len(times) / sum(times)
        sample = f"batch {b} of the {args.batch_logreg}-row workload per step, full LinearClassifier forward incl. sharing and sigmoid; dot over {threads} threads, other host ops single-threaded"
        records["logreg"] = {
            "impl": "reference", "metric": "logreg_inference_rows_per_s", "value": v, "unit": "rows/s", "steps": len(times),
            "warmup": sub_warm, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "dtype": "u128",
            "config": predictor_config(names["logreg"], world),
            "cpu_baseline": {"value": v, "unit": "rows/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        log(json.dumps(records["logreg"]))
    if args.workload in ("mlp", "logreg"):
        line = records[args.workload]
        line.update({"n_gpus": args.gpus, "scaling": "strong", "vs_baseline": None, "data": "synthetic", "gpu_launches": 0})
        print(json.dumps(line), flush=True)
        return

    size = max(args.ref_size, 1024)
    for _ in range(min(args.warmup, 1)):
        cpu_matmul(min(size, 128), threads)
    times = []
    for _ in range(args.steps):
        _, dt = cpu_matmul(size, threads)
        times.append(dt)
    value = ring_ops_per_step(size, size, size) * len(times) / sum(times) / 1e9
    sample = (f"replicated Ring128 dot {size}x{size}x{size} + TruncPr(40) per step (bounded sample of the {args.size}^3 "
              f"workload; 9 host GEMMs like the reference, OpenMP over rows on {threads} threads)")
    line = {
        "impl": "reference", "metric": "replicated_ring128_matmul_gops", "value": value, "unit": "GOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u128", "data": "synthetic",
        "config": matmul_config(names["matmul"]), "sample": sample, "note": PORT_NOTE,
        "cpu_baseline": {"value": value, "unit": "GOP/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "GOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "workloads": records,
    }
    print(json.dumps(line), flush=True)


# ---- GPU legs ----------------------------------------------------------------------------------------------------------
class Env:
    def __init__(self, args):
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        self.peaks, self.peak_src = measured_peaks()
        self.sampler = None
        self.int8_peak = None     # measured TOP/s of a library int8 GEMM (peak probe), rank 0
        self.prf_peak = None      # measured G AES blocks/s of the library's pure PRF kernel

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()


def kernel_table(trace, steps):
    """Per-launcher rows from the library trace, sorted by time; rates from the algorithmic work each launcher states."""
    total = sum(v["ms"] for v in trace.values()) or 1.0
    rows = []
    for name, v in sorted(trace.items(), key=lambda kv: -kv[1]["ms"]):
        s = v["ms"] * 1e-3
        rows.append({"kernel": name, "launches_per_step": v["n"] / steps, "ms_per_step": v["ms"] / steps,
                     "share": v["ms"] / total,
                     "GBps": (v["bytes"] / s / 1e9) if s > 0 and v["bytes"] else None,
                     "G_aes_blocks_per_s": (v["aes_blocks"] / s / 1e9) if s > 0 and v["aes_blocks"] else None,
                     "int8_TOPs": (v["tensor_ops"] / s / 1e12) if s > 0 and v["tensor_ops"] else None})
    return rows, total / steps


def roofline_from(env, rows, traced_ms_per_step, traffic=None):
    """`roofline` of the dominant launcher.  bound: "tensor" for the limb GEMM, else "hbm" (algorithmic bytes / time
    against the measured copy bandwidth); PRF-consuming kernels also carry their AES-block rate against the measured
    PRF peak, which is what actually bounds them."""
    top = rows[0]
    hbm = env.peaks["hbm_gbs"]
    if top["int8_TOPs"]:
        peak = env.int8_peak or 2.0 * env.peaks.get("bf16_tflops_sustained", env.peaks["bf16_tflops"])
        src = ("int8 GEMM peak probe measured in this run" if env.int8_peak else
               f"2 x bf16_tflops_sustained of {env.peak_src} (no int8 figure measured)")
        r = {"kernel": top["kernel"], "bound": "tensor", "achieved": top["int8_TOPs"], "peak": peak, "unit": "TOP/s (int8)",
             "frac": top["int8_TOPs"] / peak, "traffic": traffic, "peak_source": src}
    else:
        r = {"kernel": top["kernel"], "bound": "hbm", "achieved": top["GBps"], "peak": hbm, "unit": "GB/s",
             "frac": (top["GBps"] / hbm) if top["GBps"] else None, "traffic": traffic,
             "peak_source": f"hbm_gbs of {env.peak_src}"}
        if top["G_aes_blocks_per_s"]:
            r["prf"] = {"achieved": top["G_aes_blocks_per_s"], "peak": env.prf_peak, "unit": "G AES blocks/s",
                        "frac": (top["G_aes_blocks_per_s"] / env.prf_peak) if env.prf_peak else None,
                        "note": "the kernel expands PRF streams; the AES rate, not HBM, bounds it"}
    r["avg_launch_ms"] = top["ms_per_step"] / max(top["launches_per_step"], 1e-9)
    r["launches_per_step"] = top["launches_per_step"]
    r["share_of_step"] = top["share"]
    r["traced_ms_per_step"] = traced_ms_per_step
    return r


def trim(rows, frac=0.95, cap=12):
    out, acc = [], 0.0
    for r in rows:
        out.append({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()})
        acc += r["share"]
        if acc >= frac or len(out) >= cap:
            break
    return out


def probe_peaks(env):
    """Rank 0, before the workloads: (1) the library's pure PRF kernel (SampleSeeded of 2^26 Ring128 elements, 1 GB of
    keystream) = the measured PRF peak the PRF-consuming kernels are compared with; (2) an int8 tensor GEMM of a
    library (torch._int_mm -> cuBLASLt, 8192^3), used ONLY as the measured int8 peak for the limb GEMM's fraction."""
    import moose_b200 as mb
    sess = mb.Session(env.local_rank, f"{SESSION_ID}-probe", ENTROPY_KEY)
    n = 1 << 26
    for _ in range(2):
        t = sess.sample_uniform_seeded(mb.RING128, (n,), bytes(16))
    best = None
    for _ in range(5):
        sess.timer_start()
        t = sess.sample_uniform_seeded(mb.RING128, (n,), bytes(16))
        ms = sess.timer_stop()
        best = ms if best is None else min(best, ms)
    del t
    env.prf_peak = n / (best * 1e-3) / 1e9
    sess.close()
    try:
        import torch
        dev = torch.device("cuda", env.local_rank)
        m = 8192
        a = torch.randint(-128, 127, (m, m), dtype=torch.int8, device=dev)
        b = torch.randint(-128, 127, (m, m), dtype=torch.int8, device=dev)
        for _ in range(3):
            torch._int_mm(a, b)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record()
        for _ in range(reps):
            torch._int_mm(a, b)
        e1.record()
        torch.cuda.synchronize(dev)
        env.int8_peak = 2.0 * m ** 3 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12
        del a, b
        torch.cuda.empty_cache()
    except Exception as e:  # no int8 GEMM in this torch build: the 2 x bf16 proxy stays
        log(f"int8 peak probe unavailable: {e!r}")


def pipelined(sess, loader, drainer, steps, load, compute, drain):
    """Per step: the H2D of step i+1 (loader context) and the output path of step i-1 (drainer context) overlap the
    kernels of step i; contexts are ordered with mb_ctx_wait."""
    ins, keep = load(), None
    for i in range(steps):
        sess.wait(loader)                              # operands of step i are on the device
        nxt = load() if i + 1 < steps else None        # prefetch: overlaps the kernels below
        r = compute(ins)
        r.sess = drainer                               # freed on the drainer's stream, after its copy
        drainer.wait(sess)
        drain(i, r)
        loader.wait(sess)                              # the inputs are freed on the loader's stream after their last use
        keep, ins = (ins, r), nxt                      # step i's tensors live until step i+1 has been enqueued
        del r
    del keep
    drainer.sync()
    loader.sync()
    sess.sync()


def run_matmul(env):
    import moose_b200 as mb
    args, rank, world, lr = env.args, env.rank, env.world, env.local_rank
    M = N = K = args.size
    sess = mb.Session(lr, f"{SESSION_ID}-{rank}", ENTROPY_KEY)
    rng = np.random.default_rng(rank)
    hA = sess.pinned_empty((M, K, 2), np.uint64)
    hB = sess.pinned_empty((K, N, 2), np.uint64)
    hA[...] = fixed128(rng.uniform(-1, 1, (M, K)))
    hB[...] = fixed128(rng.uniform(-1, 1, (K, N)))
    x = sess.share(sess.from_host(mb.RING128, hA), 0)
    y = sess.share(sess.from_host(mb.RING128, hB), 1)
    sess.sync()

    def step():
        return sess.fixed_dot(FRAC, x, y)     # DotOp::rep_rep_kernel + TruncPrOp::rep_kernel, one call

    for _ in range(args.warmup):
        out = step()
    sess.sync()
    # ---- kernel-only timed region (inputs resident in HBM) ----
    sess.profile_enable(True)
    sess.profile_read()
    env.barrier()
    sess.sync()
    env.sampler.mark()
    launches0 = sess.launch_count
    sess.timer_start()
    for _ in range(args.steps):
        out = step()
    ms = sess.timer_stop()
    sess.sync()
    env.barrier()
    launches = sess.launch_count - launches0
    clocks = env.sampler.snapshot()
    gemm_ms, gemm_launches = sess.profile_read()
    sess.profile_enable(False)
    # ---- per-kernel trace pass (same steps) ----
    sess.trace_enable(True)
    for _ in range(args.steps):
        out = step()
    rows, traced = kernel_table(sess.trace_read(), args.steps)
    sess.trace_enable(False)
    del out

    # ---- end-to-end timed region (host buffers in, host buffer out) ----
    loader = mb.Session(lr, f"{SESSION_ID}-{rank}-loader", ENTROPY_KEY)
    drainer = mb.Session(lr, f"{SESSION_ID}-{rank}-drainer", ENTROPY_KEY)
    hOuts = [sess.pinned_empty((M, N, 2), np.uint64) for _ in range(2)]

    def load():
        return loader.from_host_async(mb.RING128, hA), loader.from_host_async(mb.RING128, hB)

    def compute(ins):
        a, b = ins
        z = sess.fixed_dot(FRAC, sess.share(a, 0), sess.share(b, 1))
        return sess.reveal(z, 0)

    def drain(i, r):
        drainer.to_host_async(r, hOuts[i % 2])

    pipelined(sess, loader, drainer, 2, load, compute, drain)
    env.barrier()
    t0 = time.perf_counter()
    pipelined(sess, loader, drainer, args.steps, load, compute, drain)
    ms_e2e = 1e3 * (time.perf_counter() - t0)
    env.barrier()
    hOut = hOuts[(args.steps - 1) % 2]
    ms, ms_e2e = reduce_max(env.dist, [ms, ms_e2e], lr)
    launches = reduce_sum_int(env.dist, launches, lr)

    # spot check: the revealed product decodes to A @ B (plaintext) within the TruncPr error
    a0 = hA[:4, :, 0].astype(np.int64).astype(np.float64) / 2 ** FRAC
    b0 = hB[:, :4, 0].astype(np.int64).astype(np.float64) / 2 ** FRAC
    got = hOut[:4, :4, 0].astype(np.int64).astype(np.float64) / 2 ** FRAC
    ok = bool(np.allclose(got, a0 @ b0, atol=K * 2.0 ** -FRAC * 4 + 2.0 ** -38))
    h2d, d2h = int(hA.nbytes + hB.nbytes), int(hOut.nbytes)
    loader.close()
    drainer.close()
    del x, y
    sess.close()
    if rank != 0:
        return None
    ops = ring_ops_per_step(M, K, N)
    name = workload_names(args.batch_mlp, args.batch_logreg, args.size)["matmul"]
    # the limb GEMM's own (start, stop) events inside the timed region
    int8_ops = 3 * 2.0 * M * N * (2 * K) * 136
    avg_ms = gemm_ms / max(gemm_launches, 1)
    achieved = int8_ops / (avg_ms * 1e-3) / 1e12 if gemm_launches else None
    peak = env.int8_peak or 2.0 * env.peaks.get("bf16_tflops_sustained", env.peaks["bf16_tflops"])
    line = {
        "metric": "replicated_ring128_matmul_gops", "value": aggregate_value(ops, args.steps, world, ms) / 1e9, "unit": "GOP/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u128", "data": "synthetic",
        "config": matmul_config(name), "result_check": "ok" if ok else "MISMATCH",
        "clocks": clocks,
        "e2e": {"value": aggregate_value(ops, args.steps, world, ms_e2e) / 1e9, "unit": "GOP/s", "ms_per_step": ms_e2e / args.steps,
                "pipelining": "per step: H2D of both operands and D2H of the result, pinned host memory; copies of "
                              "steps i+1 / i-1 overlap the kernels of step i (loader / drainer contexts)",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches),
        "roofline": {"kernel": "limb_gemm_kernel_2cta", "bound": "tensor", "achieved": achieved, "peak": peak,
                     "unit": "TOP/s (int8)", "frac": (achieved / peak) if achieved else None,
                     # dram__bytes_read.sum + dram__bytes_write.sum of one launch (ncu --set full, profiles/); the
                     # algorithmic operand + partial bytes of the launch are 2.4 GB + 4.0 GB
                     "traffic": GEMM_TRAFFIC_BYTES if (M, K, N) == (4096, 4096, 4096) else None,
                     "avg_launch_ms": avg_ms, "launches_timed": gemm_launches, "share_of_step": gemm_ms / ms if ms else None,
                     "peak_source": ("int8 GEMM peak probe measured in this run (torch._int_mm 8192^3, library kernel, probe only)"
                                     if env.int8_peak else f"2 x bf16_tflops_sustained of {env.peak_src} (no int8 figure measured)"),
                     "peak_2x_bf16_sustained": 2.0 * env.peaks.get("bf16_tflops_sustained", env.peaks["bf16_tflops"])},
        "kernels": trim(rows),
        "peaks": {"hbm_gbs": env.peaks["hbm_gbs"], "int8_tops_probe": env.int8_peak, "prf_g_blocks_per_s": env.prf_peak,
                  "source": env.peak_src},
    }
    return line


GEMM_TRAFFIC_BYTES = 43.43e9   # profiles/r2_matmul_step_ncu_full_raw.csv
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernels that dominate the predictor steps, from the
# `ncu --set full` captures of one step at the BASELINE batch sizes on one GPU (profiles/r2_ncu_kernel_table.md)
TRAFFIC_BYTES = {("mlp", "limb_gemm"): 9.06e9, ("mlp", "maximum_select"): 2.40e9, ("mlp", "adder_level"): 0.74e9,
                 ("logreg", "rep_dot_skinny_deferred"): 8.40e9, ("logreg", "inject_sum"): 0.129e9, ("logreg", "adder_level"): 0.27e9}


def run_predictor(env, kind):
    """mlp / logreg: one batch split by rows over the ranks; kernel-only value, trace pass, pipelined e2e with gather."""
    import moose_b200 as mb
    from moose_b200.predictors import LinearPredictor, MLPPredictor
    args, rank, world, lr = env.args, env.rank, env.world, env.local_rank
    if kind == "mlp":
        B, cols_in, cols_out, ring, frac, seed = args.batch_mlp, MLP_DIMS[0], MLP_DIMS[-1], mb.RING64, MLP_FRAC, 3
        lo_hi = (0.0, 1.0)
    else:
        B, cols_in, cols_out, ring, frac, seed = args.batch_logreg, LOGREG_FEATS, 2, mb.RING128, LOGREG_FRAC, 1
        lo_hi = (-1.0, 1.0)
    rows_of = [b - a for a, b in (shard_rows(B, world, g) for g in range(world))]
    rows_local = rows_of[rank]
    sess = mb.Session(lr, f"{SESSION_ID}-{kind}-{rank}", ENTROPY_KEY)
    if kind == "mlp":
        ws, bs = mlp_model()
        pred = MLPPredictor(sess, ws, bs, ring, frac, "maximum", integral=MLP_INT, bias_via="share")
    else:
        coeffs, intercepts = logreg_model()
        pred = LinearPredictor(sess, coeffs, intercepts, ring, frac, LOGREG_INT, "sigmoid")
    hX = sess.pinned_empty((rows_local, cols_in), np.float64)
    fill_uniform(hX, seed * 1000 + rank, *lo_hi)
    head = hX[:64].copy()

    xf = pred.encode(hX)          # device-resident fixed-point batch
    sess.sync()

    def step():
        return sess.reveal(pred.forward_shared(xf), 0)

    for _ in range(args.warmup):
        out = step()
    sess.sync()
    env.barrier()
    sess.sync()
    env.sampler.mark()
    l0 = sess.launch_count
    sess.timer_start()
    for _ in range(args.steps):
        out = step()
    ms = sess.timer_stop()
    sess.sync()
    env.barrier()
    launches = sess.launch_count - l0
    clocks = env.sampler.snapshot()
    got = sess.fixedpoint_decode(frac, out).to_host()[:64]
    # trace pass
    sess.trace_enable(True)
    for _ in range(args.steps):
        out = step()
    rows, traced = kernel_table(sess.trace_read(), args.steps)
    sess.trace_enable(False)
    del out, xf

    # ---- end to end: host float64 batch -> decoded outputs of all ranks on rank 0's host ----
    loader = mb.Session(lr, f"{SESSION_ID}-{kind}-{rank}-loader", ENTROPY_KEY)
    drainer = mb.Session(lr, f"{SESSION_ID}-{kind}-{rank}-drainer", ENTROPY_KEY)
    gather = OutputGather(env.dist, drainer, lr, rows_of, cols_out, world, rank)

    def load():
        return loader.from_host_async(mb.FLOAT64, hX)

    def compute(xin):
        x_fixed = sess.fixedpoint_encode(ring, frac, xin)
        return sess.fixedpoint_decode(frac, sess.reveal(pred.forward_shared(x_fixed), 0))

    pipelined(sess, loader, drainer, 2, load, compute, gather.step)
    env.barrier()
    t0 = time.perf_counter()
    pipelined(sess, loader, drainer, args.steps, load, compute, gather.step)
    if env.dist is not None:
        import torch
        torch.cuda.synchronize(lr)
    ms_e2e = 1e3 * (time.perf_counter() - t0)
    env.barrier()
    final = gather.result(args.steps - 1)
    if final is not None:
        final = np.array(final[:64], copy=True), final.shape, bool(np.isfinite(final).all())
    gather.close()
    ms, ms_e2e = reduce_max(env.dist, [ms, ms_e2e], lr)
    launches = reduce_sum_int(env.dist, launches, lr)
    if kind == "mlp":
        ref = mlp_float(head, ws, bs)
        tol = 5e-3
    else:
        ref = logreg_float(head, coeffs, intercepts)
        tol = 2e-2   # the reference's classifier tests check 2 decimals (Goldschmidt at fixed(24,40))
    err = float(np.abs(got - ref).max())
    ok = err < tol
    if final is not None:
        err_e2e = float(np.abs(final[0] - ref).max())   # rank 0's rows come first
        ok = ok and err_e2e < tol and final[1] == (B, cols_out) and final[2]
    loader.close()
    drainer.close()
    del pred
    sess.close()
    if rank != 0:
        return None
    names = workload_names(args.batch_mlp, args.batch_logreg, args.size)
    rec = {
        "metric": f"{kind}_inference_rows_per_s", "value": B * args.steps / (ms * 1e-3), "unit": "rows/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "u64" if kind == "mlp" else "u128", "data": "synthetic",
        "config": predictor_config(names[kind], world),
        "result_check": "ok" if ok else "MISMATCH", "max_abs_err_vs_float": err,
        "clocks": clocks,
        "e2e": {"value": B * args.steps / (ms_e2e * 1e-3), "unit": "rows/s", "ms_per_step": ms_e2e / args.steps,
                "pipelining": "per step and rank: H2D of the float64 rows from pinned memory, encode, share, forward, reveal, "
                              "decode; NCCL gather of the decoded outputs to rank 0 and its D2H; copies overlap the kernels of "
                              "the neighbouring steps (loader / drainer contexts)",
                "h2d_bytes_per_step": int(B * cols_in * 8), "d2h_bytes_per_step": int(B * cols_out * 8)},
        "gpu_launches": int(launches),
        "roofline": roofline_from(env, rows, traced,
                                  TRAFFIC_BYTES.get((kind, rows[0]["kernel"])) if (world == 1 and B == (65536 if kind == "mlp" else 1000000)) else None),
        "kernels": trim(rows),
    }
    return rec


def run_party(env):
    """BASELINE configs[4]: the MLP with alice / bob / carole on GPUs 0 / 1 / 2 (ranks 0..2 of a job with >= 4 ranks; the
    other ranks wait), party-sliced fused kernels, reshares and openings stored into the receivers' HBM over NVLink
    (csrc/peer_transport.cu).  Timed by the host clock around `repeat` full inferences bracketed by barriers: the three
    streams wait for each other's messages, so no single device clock spans the run."""
    import argparse as _ap
    from moose_b200 import party as party_mod
    args = env.args
    group = env.dist.new_group([0, 1, 2])           # every rank of the job takes part in creating it
    rec = None
    if env.rank < 3:
        pa = _ap.Namespace(session_id=f"{SESSION_ID}-party", native_nccl=False, peer_memory=True, lanes=1, ring_mib=1024,
                           dims=list(MLP_DIMS), mlp=args.batch_mlp, repeat=max(args.steps, 1))
        rec = party_mod.run_fused_mlp(pa, group=group)
    env.barrier()
    if env.rank != 0 or rec is None:
        return None
    ms = 1e3 * rec["seconds_per_run"]
    return {"metric": "mlp_inference_rows_per_s", "value": rec["rows_per_s"], "unit": "rows/s", "n_gpus": 3, "ms": ms,
            "ms_per_step": ms, "rows_per_s": rec["rows_per_s"], "steps": max(args.steps, 1), "warmup": 1,
            "nvlink_bytes": rec["payload_bytes_per_run"], "transport": rec["transport"],
            "config": {"workload": rec["workload"], "parties": "alice / bob / carole on GPUs 0 / 1 / 2, one process each"},
            "max_abs_err_first_rows": rec["max_abs_err_first_rows"],
            "result_check": "ok" if (rec["max_abs_err_first_rows"] is not None and rec["max_abs_err_first_rows"] < 5e-3) else "MISMATCH",
            "timing": "host clock around the inferences, barriers on both sides (three dependent streams)",
            "per_party": rec["per_party"]}


def add_cpu_baselines(env, line, records):
    """Rank 0 at N = 1: the port on bounded samples, single-threaded host ops like the reference."""
    args = env.args
    threads = host_threads()
    if line is not None:
        v, dt = cpu_matmul(args.ref_size, 1)
        line["cpu_baseline"] = {
            "value": v, "unit": "GOP/s", "cores": 1, "kind": "port",
            "sample": f"replicated Ring128 dot {args.ref_size}^3 + TruncPr(40), {dt:.1f} s; {PORT_NOTE}; box has {threads} usable cores"}
    if "mlp" in records:
        b = args.cpu_batch_mlp
        v, dt = cpu_mlp(b, 1)
        v3, dt3 = cpu_mlp(b, 3)
        records["mlp"]["cpu_baseline"] = {
            "value": v, "unit": "rows/s", "cores": 1, "kind": "port",
            "sample": f"batch {b} of the {args.batch_mlp}-row workload, full forward, {dt:.1f} s; {PORT_NOTE}",
            "three_threads": {"value": v3, "cores": 3, "seconds": dt3, "note": "dot over 3 threads (one per party's GEMMs)"}}
    if "logreg" in records:
        b = args.cpu_batch_logreg
        v, dt = cpu_logreg(b, 1)
        v3, dt3 = cpu_logreg(b, 3)
        records["logreg"]["cpu_baseline"] = {
            "value": v, "unit": "rows/s", "cores": 1, "kind": "port",
            "sample": f"batch {b} of the {args.batch_logreg}-row workload, full forward incl. sigmoid, {dt:.1f} s; {PORT_NOTE}",
            "three_threads": {"value": v3, "cores": 3, "seconds": dt3, "note": "dot over 3 threads"}}


def run_prf(args):
    """PRF microbenchmark (the ceiling of the PRF-consuming kernels): SampleSeeded of a 4 GB Ring128 tensor.
    Stated bound: the T-table formulation does 148 shared-memory lookups per AES block (4 + 8 x 16 + 16) and an SM retires
    one 32-lane LDS per clock, so peak = SMs x 32 x f_SM / 148 blocks/s at the SM clock sampled under load."""
    import moose_b200 as mb
    sess = mb.Session(0, f"{SESSION_ID}-prf", ENTROPY_KEY)
    n = 1 << 28
    seed = bytes(range(16))
    for _ in range(args.warmup):
        t = sess.sample_uniform_seeded(mb.RING128, (n,), seed)
    sess.sync()
    sampler = ClockSampler(0)
    sampler.start()
    sess.timer_start()
    for _ in range(args.steps):
        t = sess.sample_uniform_seeded(mb.RING128, (n,), seed)
    ms = sess.timer_stop()
    clocks = sampler.stop()
    blocks_per_s = n * args.steps / (ms * 1e-3)
    f_hz = (clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0) * 1e6
    peak = 148 * 32 * f_hz / 148.0
    first = sess.sample_uniform_seeded(mb.RING128, (1,), bytes(16)).to_ints().tolist()
    print(json.dumps({
        "metric": "prf_aes128_ctr_blocks_per_s", "value": blocks_per_s / 1e9, "unit": "G blocks/s", "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "sample_seeded_ring128_2^28_elements (4 GB of keystream per step)",
                   "golden_first_element_ok": first == [78655730786844307471556614669614075016]},
        "clocks": clocks, "gpu_launches": args.steps,
        "roofline": {"kernel": "sample_kernel", "bound": "shared-memory pipe (LDS)", "achieved": blocks_per_s / 1e9,
                     "peak": peak / 1e9, "unit": "G blocks/s", "frac": blocks_per_s / peak,
                     "keystream_GBps": blocks_per_s * 16 / 1e9,
                     "peak_source": "148 SMs x 32 lanes x sampled SM clock / 148 lookups per block"}}), flush=True)
    sess.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="all", choices=["all", "matmul", "mlp", "logreg", "prf", "party"],
                    help="all = matmul (the top-level fields) + mlp + logreg records in one line")
    ap.add_argument("--size", type=int, default=4096, help="matmul M=N=K (default: the BASELINE config, 4096)")
    ap.add_argument("--batch-mlp", type=int, default=65536)
    ap.add_argument("--batch-logreg", type=int, default=1000000)
    ap.add_argument("--batch", type=int, default=None, help="overrides the batch of --workload mlp / logreg")
    ap.add_argument("--ref-size", type=int, default=640, help="matmul sample size of the CPU baseline (>= 1024 in --impl reference)")
    ap.add_argument("--cpu-batch-mlp", type=int, default=256)
    ap.add_argument("--cpu-batch-logreg", type=int, default=1024)
    ap.add_argument("--ref-batch-mlp", type=int, default=512)
    ap.add_argument("--ref-batch-logreg", type=int, default=2048)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-peak-probe", action="store_true")
    args = ap.parse_args()
    if args.batch is not None:
        args.batch_mlp = args.batch_logreg = args.batch
    if args.workload == "prf":
        return run_prf(args)

    env = Env(args)
    if args.impl == "reference":
        return run_reference(args, env.rank, env.world)

    env.dist = dist_setup(env.world, env.local_rank)
    env.sampler = ClockSampler(env.local_rank)
    env.sampler.start()
    if not args.no_peak_probe:
        probe_peaks(env)
    line, records = None, {}
    if args.workload in ("all", "matmul"):
        line = run_matmul(env)
    for kind in ("mlp", "logreg"):
        if args.workload in ("all", kind):
            rec = run_predictor(env, kind)
            if rec is not None:
                records[kind] = rec
    party = None
    if env.world >= 4 and args.workload in ("all", "party"):
        party = run_party(env)
    env.sampler.stop()
    if env.rank == 0:
        if party is not None and line is not None:
            line["party_per_gpu"] = party
        elif party is not None:
            line = party
        if env.world == 1 and not args.no_cpu_baseline:
            add_cpu_baselines(env, line, records)
        for rec in records.values():
            log(json.dumps(rec))
        if line is None:                      # a single predictor workload: its record is the line
            line = records[args.workload]
        else:
            line["workloads"] = records
        print(json.dumps(line), flush=True)
    if env.dist is not None:
        env.dist.destroy_process_group()


if __name__ == "__main__":
    main()
