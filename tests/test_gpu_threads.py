"""Concurrent callers (include/moose_b200.h: "thread safe"; the reference runs every op of a session as its own
tokio task, moose/src/execution/asynchronous.rs:453-531).

  * four host threads issue 50 mixed calls each (tensor-pipe dot, IMAD dot, mul, seeded sampling) on ONE context;
    every result is compared bit for bit with the oracle.  The tensor-pipe GEMM keeps limb planes and partial sums
    in per-context scratch, so split -> GEMM -> recombine of one call must reach the stream as a unit.
  * two contexts on one GPU run the 2-CTA tensor-pipe GEMM at the same time (the in-process three-party mode and the
    loader / compute / drainer contexts of bench.py share a device like this): both results exact, no hang.
  * concurrent first uses of the replicated setup on one context see ONE setup (zero shares cancel).
"""
import threading

import numpy as np
import pytest

import moose_b200 as mb
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY, rand_ring

pytestmark = pytest.mark.gpu
DT = {64: mb.RING64, 128: mb.RING128}


def _run_threads(fns):
    errs = []

    def wrap(f):
        try:
            f()
        except BaseException as e:  # noqa: BLE001  (reported in the main thread)
            errs.append(e)

    ts = [threading.Thread(target=wrap, args=(f,)) for f in fns]
    for t in ts:
        t.start()
    for t in ts:
        t.join(timeout=300)
        assert not t.is_alive(), "a worker thread hung"
    if errs:
        raise errs[0]


def test_four_threads_mixed_calls_on_one_context():
    sess = mb.Session(0, "b200-threads", ENTROPY_KEY)
    sess.set_gemm_mode(0)   # automatic: large dots take the tensor pipe, small ones the IMAD kernel
    results = [[] for _ in range(4)]

    def worker(tid):
        rng = np.random.default_rng(100 + tid)
        for i in range(50):
            kind = (i + tid) % 4
            bits = 128 if (i // 4 + tid) % 2 else 64
            if kind == 0:      # tensor-pipe GEMM (scratch planes), ragged K on some calls
                M, K, N = 512, 512 + 64 * (i % 3) + (tid == 1) * 3, 512   # M*N*K >= 2^27: automatic mode picks the tensor pipe
                a, b = rand_ring(rng, bits, (M, K)), rand_ring(rng, bits, (K, N))
                got = sess.dot(sess.from_host(DT[bits], a), sess.from_host(DT[bits], b)).to_host()
                want = mo.dot(mo.Ring(bits, a), mo.Ring(bits, b)).data
            elif kind == 1:    # small dot (IMAD)
                a, b = rand_ring(rng, bits, (17, 33)), rand_ring(rng, bits, (33, 5))
                got = sess.dot(sess.from_host(DT[bits], a), sess.from_host(DT[bits], b)).to_host()
                want = mo.dot(mo.Ring(bits, a), mo.Ring(bits, b)).data
            elif kind == 2:
                a, b = rand_ring(rng, bits, (1000 + i,)), rand_ring(rng, bits, (1000 + i,))
                got = sess.mul(sess.from_host(DT[bits], a), sess.from_host(DT[bits], b)).to_host()
                want = mo.mul(mo.Ring(bits, a), mo.Ring(bits, b)).data
            else:
                seed = bytes([tid, i] + [7] * 14)
                got = sess.sample_uniform_seeded(DT[bits], (777 + i,), seed).to_host()
                want = mo.sample_uniform_seeded(bits, (777 + i,), seed).data
            results[tid].append(bool(np.array_equal(got, want)))

    try:
        _run_threads([lambda t=t: worker(t) for t in range(4)])
    finally:
        sess.close()
    for tid in range(4):
        assert len(results[tid]) == 50 and all(results[tid]), (tid, results[tid])


def test_two_contexts_run_the_pair_gemm_concurrently():
    rng = np.random.default_rng(5)
    M, K, N = 1024, 1024, 512
    cases = []
    for c in range(2):
        a, b = rand_ring(rng, 128, (M, K)), rand_ring(rng, 128, (K, N))
        cases.append((a, b, mo.dot(mo.Ring(128, a), mo.Ring(128, b)).data))
    sessions = [mb.Session(0, f"b200-pair-{c}", ENTROPY_KEY) for c in range(2)]
    ok = [[], []]

    def worker(c):
        s = sessions[c]
        s.set_gemm_mode(2)
        a, b, want = cases[c]
        A, B = s.from_host(mb.RING128, a), s.from_host(mb.RING128, b)
        for _ in range(6):
            ok[c].append(bool(np.array_equal(s.dot(A, B).to_host(), want)))

    try:
        _run_threads([lambda c=c: worker(c) for c in range(2)])
    finally:
        for s in sessions:
            s.close()
    assert ok == [[True] * 6, [True] * 6]


def test_concurrent_first_use_of_the_replicated_setup():
    sess = mb.Session(0, "b200-setup-race", ENTROPY_KEY)
    rng = np.random.default_rng(9)
    vals = [rand_ring(rng, 64, (4096,)) for _ in range(4)]
    out = [None] * 4

    def worker(t):
        x = sess.share(sess.from_host(mb.RING64, vals[t]), t % 3)
        y = sess.rep_mul(x, x)           # needs the setup keys (zero share)
        out[t] = sess.reveal(y).to_host()

    try:
        _run_threads([lambda t=t: worker(t) for t in range(4)])
    finally:
        sess.close()
    for t in range(4):
        assert np.array_equal(out[t], vals[t] * vals[t])   # uint64 wraps like the ring
