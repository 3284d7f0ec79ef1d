"""IMAD vs tensor-pipe ring GEMM: device time of mb_dot in gemm mode 1 (64-bit IMAD kernels) and mode 2 (tcgen05 int8 limb
GEMM) over cubic and tall shapes -- the evidence behind k_gemm_limb_preferred (csrc/gemm_limb.cu: N >= 64 and
M*N*K >= 2^27).  Run on a B200:  python tools/gemm_crossover.py > gpurun_out/gemm_crossover.txt"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import moose_b200 as mb  # noqa: E402


def timed(sess, fn, reps=5):
    fn()
    sess.sync()
    sess.timer_start()
    for _ in range(reps):
        fn()
    return sess.timer_stop() / reps


def main():
    sess = mb.Session(0, "crossover", bytes(range(16)))
    print("| dtype | M | K | N | M*N*K | IMAD ms | tensor pipe ms | faster | automatic mode picks |")
    print("|---|---|---|---|---|---|---|---|---|")
    shapes = [(128, 128, 128), (256, 256, 256), (384, 384, 384), (512, 512, 512), (768, 768, 768), (1024, 1024, 1024),
              (2048, 2048, 2048), (4096, 256, 128), (4096, 512, 64), (8192, 512, 512), (65536, 784, 512), (1024, 4096, 64)]
    for dt, name in ((mb.RING64, "ring64"), (mb.RING128, "ring128")):
        for (M, K, N) in shapes:
            a = sess.sample_uniform_seeded(dt, (M, K), bytes(16))
            b = sess.sample_uniform_seeded(dt, (K, N), bytes([1] * 16))
            t = {}
            for mode in (1, 2):
                sess.set_gemm_mode(mode)
                try:
                    t[mode] = timed(sess, lambda: sess.dot(a, b))
                except mb.MooseB200Error:
                    t[mode] = None
            sess.set_gemm_mode(0)
            auto = "tensor pipe" if (N >= 64 and M * N * K >= (1 << 27)) else "IMAD"
            fast = "tensor pipe" if (t[2] is not None and t[2] < t[1]) else "IMAD"
            print(f"| {name} | {M} | {K} | {N} | 2^{np.log2(M * N * K):.1f} | {t[1]:.3f} | "
                  f"{'n/a' if t[2] is None else format(t[2], '.3f')} | {fast} | {auto}{'' if fast == auto else ' (!)'} |")
    sess.close()


if __name__ == "__main__":
    main()
