"""One step of a bench workload between cudaProfilerStart / cudaProfilerStop, for Nsight Compute:

    ncu --profile-from-start off --set full --clock-control none --import-source on -o gpurun_out/<name> \
        python tools/profile_step.py --workload mlp|logreg|matmul [--batch B]
    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file ... (launch list)

The step is bench.py's: same inputs, same calls (python bench.py is what produces the numbers; nothing printed here
is a measurement)."""
import argparse
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import moose_b200 as mb  # noqa: E402
from moose_b200.predictors import LinearPredictor, MLPPredictor  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="mlp", choices=["mlp", "logreg", "matmul"])
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--size", type=int, default=4096)
    args = ap.parse_args()
    cudart = ctypes.CDLL("libcudart.so")
    sess = mb.Session(0, f"{bench.SESSION_ID}-profile", bench.ENTROPY_KEY)
    if args.workload == "matmul":
        n = args.size
        rng = np.random.default_rng(0)
        x = sess.share(sess.from_host(mb.RING128, bench.fixed128(rng.uniform(-1, 1, (n, n)))), 0)
        y = sess.share(sess.from_host(mb.RING128, bench.fixed128(rng.uniform(-1, 1, (n, n)))), 1)
        step = lambda: sess.fixed_dot(bench.FRAC, x, y)
    elif args.workload == "mlp":
        B = args.batch or 65536
        ws, bs = bench.mlp_model()
        pred = MLPPredictor(sess, ws, bs, mb.RING64, bench.MLP_FRAC, "maximum", integral=bench.MLP_INT, bias_via="share")
        xf = pred.encode(np.random.default_rng(3).uniform(0, 1, (B, bench.MLP_DIMS[0])))
        step = lambda: sess.reveal(pred.forward_shared(xf), 0)
    else:
        B = args.batch or 1000000
        coeffs, intercepts = bench.logreg_model()
        pred = LinearPredictor(sess, coeffs, intercepts, mb.RING128, bench.LOGREG_FRAC, bench.LOGREG_INT, "sigmoid")
        hx = np.empty((B, bench.LOGREG_FEATS), dtype=np.float64)
        bench.fill_uniform(hx, 1000, -1.0, 1.0)
        xf = pred.encode(hx)
        del hx
        step = lambda: sess.reveal(pred.forward_shared(xf), 0)
    for _ in range(2):
        out = step()
    sess.sync()
    cudart.cudaProfilerStart()
    out = step()
    sess.sync()
    cudart.cudaProfilerStop()
    del out
    print(f"profiled one {args.workload} step, {sess.launch_count} launches in the process")
    sess.close()


if __name__ == "__main__":
    main()
