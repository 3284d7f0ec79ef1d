"""Generates tests/golden/fullsize.json: SHA-256 of every output share of the oracle at the BASELINE sizes.

    python tests/golden/make_fullsize_golden.py [c2] [c4] [c3]      (minutes per case on 8 cores, run offline)

The oracle (oracle/, the CPU restatement of the reference's kernels and protocols) is too slow to run at these sizes
inside the GPU suite, so its outputs are recorded here once and the GPU tests (tests/test_gpu_fullsize.py) compare
the library's shares against the digests: bit-exact parity at the sizes bench.py measures --

  c2  replicated Ring128 fixed(24,40) 4096^3 dot (the two-term K' = 8192 ten-pass tensor-pipe plan with the wrapping
      diagonals) and TruncPr(40)
  c4  MLP 784-512-512-10, maximum([0, z]), Ring64 fixed(14,23), batch 65536, the reference's op sequence
  c3  LinearClassifier 512 features, 2 classes, sigmoid, Ring128 fixed(24,40), batch 2^18

Inputs are numpy default_rng streams (stated below), identical on every machine; session ids and the entropy key are
part of the fixture.  Only the dot products use several threads (integer arithmetic mod 2^k is order independent).
"""
import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import moose_oracle as mo, predictors_oracle as po  # noqa: E402

OUT = os.path.join(HERE, "fullsize.json")
KEY = bytes(range(16))


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def fixed128(vals, frac=40):
    scaled = np.trunc(vals * float(2 ** frac)).astype(np.int64)
    lo = scaled.astype(np.uint64)
    hi = np.where(scaled < 0, np.uint64(0xFFFFFFFFFFFFFFFF), np.uint64(0)).astype(np.uint64)
    return np.stack([lo, hi], axis=-1)


def c2_inputs(n=4096):
    rng = np.random.default_rng(0)
    return fixed128(rng.uniform(-1, 1, (n, n))), fixed128(rng.uniform(-1, 1, (n, n)))


def c4_inputs(batch=65536):
    dims = (784, 512, 512, 10)
    ws = [np.random.default_rng(4 + i).normal(0, dims[i] ** -0.5, (dims[i], dims[i + 1])) for i in range(3)]
    bs = [np.random.default_rng(40 + i).normal(0, 0.1, dims[i + 1]) for i in range(3)]
    x = np.random.default_rng(3).uniform(0, 1, (batch, dims[0]))
    return x, ws, bs


def c3_inputs(batch=1 << 18):
    coeffs = np.random.default_rng(2).normal(0, 512 ** -0.5, (2, 512))
    intercepts = np.random.default_rng(22).normal(0, 0.1, (1, 2))
    x = np.random.default_rng(1).uniform(-1, 1, (batch, 512))
    return x, coeffs, intercepts


def make_c2():
    a, b = c2_inputs()
    sess = mo.Session("b200-fullsize-c2", KEY)
    x, y = mo.rep_share(sess, mo.Ring(128, a), 0), mo.rep_share(sess, mo.Ring(128, b), 1)
    z = mo.rep_dot(sess, x, y)
    t = mo.rep_trunc_pr(sess, 40, z)
    rows = [0, 127, 128, 2047, 4095]
    return {"session": "b200-fullsize-c2", "size": 4096, "trunc": 40,
            "dot_shares_sha256": [sha(s.data) for s in z.distinct()],
            "trunc_shares_sha256": [sha(s.data) for s in t.distinct()],
            "rows": rows,
            "dot_share0_rows_sha256": [sha(z.distinct()[0].data[r]) for r in rows],
            "revealed_sha256": sha(mo.rep_reveal(t).data)}


def make_c4():
    x, ws, bs = c4_inputs()
    sess = mo.Session("b200-fullsize-c4", KEY)
    out, dec = po.mlp(sess, x, ws, bs, bits=64, frac=23, relu_via="maximum", integral=14, bias_via="share")
    return {"session": "b200-fullsize-c4", "batch": 65536, "shares_sha256": [sha(s.data) for s in out.distinct()],
            "first_rows_share0": out.distinct()[0].data[:128].reshape(-1).tolist(),
            "last_rows_share0": out.distinct()[0].data[-128:].reshape(-1).tolist(),
            "decoded_sha256": sha(dec), "decoded_first_row": dec[0].tolist(), "decoded_last_row": dec[-1].tolist()}


def make_c3():
    x, coeffs, intercepts = c3_inputs()
    sess = mo.Session("b200-fullsize-c3", KEY)
    out, dec = po.linear(sess, x, coeffs, intercepts, bits=128, frac=40, integral=24, post_transform="sigmoid")
    return {"session": "b200-fullsize-c3", "batch": 1 << 18, "shares_sha256": [sha(s.data) for s in out.distinct()],
            "decoded_sha256": sha(dec), "decoded_first_row": dec[0].tolist(), "decoded_last_row": dec[-1].tolist()}


if __name__ == "__main__":
    which = sys.argv[1:] or ["c2", "c4", "c3"]
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    mo.DOT_THREADS = mo._lib.orc_max_threads()
    for name in which:
        t0 = time.perf_counter()
        data[name] = {"c2": make_c2, "c4": make_c4, "c3": make_c3}[name]()
        data[name]["oracle_seconds"] = round(time.perf_counter() - t0, 1)
        print(name, data[name]["oracle_seconds"], "s", flush=True)
        with open(OUT, "w") as f:
            json.dump(data, f, indent=1)
