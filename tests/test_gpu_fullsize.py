"""Bit-exact parity at the BASELINE sizes (the sizes bench.py measures), against digests of the oracle's shares
recorded offline by tests/golden/make_fullsize_golden.py (the oracle needs minutes per case at these sizes):

  c2  replicated Ring128 4096^3 dot on the tensor pipe -- the two-term K' = 8192 ten-pass plan with the wrapping
      diagonals, which no smaller case reaches -- and TruncPr(40), every share of both results
  c4  MLP 784-512-512-10 at batch 65536 (grid-stride tails of every fused kernel, 3.4e7-element bit planes)
  c3  LinearClassifier + sigmoid at batch 2^18

and the AES-CTR stream across the 2^32-block boundary (the generic round-1 path of csrc/aes.cuh), which no tensor
below 64 GB reaches, through the stream-seek entry point against the oracle's AES-CTR at the same counter.
"""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

import moose_b200 as mb
from moose_b200.predictors import LinearPredictor, MLPPredictor
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_fullsize_golden as gold  # noqa: E402  (input generators shared with the fixture script)

pytestmark = pytest.mark.gpu


def _golden(name):
    path = os.path.join(HERE, "golden", "fullsize.json")
    if not os.path.exists(path):
        pytest.fail("tests/golden/fullsize.json is missing (python tests/golden/make_fullsize_golden.py)")
    data = json.load(open(path))
    if name not in data:
        pytest.fail(f"tests/golden/fullsize.json has no '{name}' entry")
    return data[name]


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_c2_ring128_4096_dot_and_trunc_every_share():
    g = _golden("c2")
    a, b = gold.c2_inputs(g["size"])
    s = mb.Session(0, g["session"], ENTROPY_KEY)
    try:
        s.set_gemm_mode(2)   # tensor-pipe limb GEMM, CTA pairs (what bench.py's matmul runs)
        x = s.share(s.from_host(mb.RING128, a), 0)
        y = s.share(s.from_host(mb.RING128, b), 1)
        z = s.rep_dot(x, y)
        t = s.trunc_pr(g["trunc"], z)
        z0 = z.share(0).to_host()
        assert [_sha(z0[r]) for r in g["rows"]] == g["dot_share0_rows_sha256"]
        assert [_sha(z0)] + [_sha(z.share(i).to_host()) for i in (1, 2)] == g["dot_shares_sha256"]
        assert [_sha(t.share(i).to_host()) for i in range(3)] == g["trunc_shares_sha256"]
        assert _sha(s.reveal(t).to_host()) == g["revealed_sha256"]
    finally:
        s.close()


def test_c4_mlp_batch_65536_every_share():
    g = _golden("c4")
    x, ws, bs = gold.c4_inputs(g["batch"])
    s = mb.Session(0, g["session"], ENTROPY_KEY)
    try:
        pred = MLPPredictor(s, ws, bs, mb.RING64, 23, "maximum", integral=14, bias_via="share")
        out = pred.forward_shared(pred.encode(x))
        s0 = out.share(0).to_host()
        assert s0[:128].reshape(-1).tolist() == g["first_rows_share0"]
        assert s0[-128:].reshape(-1).tolist() == g["last_rows_share0"]
        assert [_sha(s0)] + [_sha(out.share(i).to_host()) for i in (1, 2)] == g["shares_sha256"]
        dec = s.fixedpoint_decode(23, s.reveal(out)).to_host()
        assert _sha(dec) == g["decoded_sha256"]
        assert dec[0].tolist() == g["decoded_first_row"] and dec[-1].tolist() == g["decoded_last_row"]
    finally:
        s.close()


def test_c3_logreg_batch_2p18_every_share():
    g = _golden("c3")
    x, coeffs, intercepts = gold.c3_inputs(g["batch"])
    s = mb.Session(0, g["session"], ENTROPY_KEY)
    try:
        pred = LinearPredictor(s, coeffs, intercepts, mb.RING128, 40, 24, "sigmoid")
        out = pred.forward_shared(pred.encode(x))
        assert [_sha(out.share(i).to_host()) for i in range(3)] == g["shares_sha256"]
        dec = s.fixedpoint_decode(40, s.reveal(out)).to_host()
        assert _sha(dec) == g["decoded_sha256"]
        assert dec[0].tolist() == g["decoded_first_row"] and dec[-1].tolist() == g["decoded_last_row"]
    finally:
        s.close()


@pytest.mark.parametrize("dtype,bits", [(mb.RING128, 128), (mb.RING64, 64), (mb.BIT, 0)])
@pytest.mark.parametrize("first_block", [0, 7, (1 << 32) - 300, 1 << 32, (1 << 40) + 5])
def test_aes_ctr_stream_seek_across_2p32_blocks(dtype, bits, first_block):
    """Vector v of the tensor = AES-128-CTR block first_block + v; blocks at and beyond 2^32 take the generic first
    round of aes_ctr_block (csrc/aes.cuh).  Oracle: the C AES-CTR at the same starting counter."""
    import ctypes
    seed = bytes(range(50, 66))
    nblocks = 1000
    s = mb.Session(0, "b200-seek", ENTROPY_KEY)
    try:
        shape = (nblocks,) if bits == 128 else (2 * nblocks,) if bits == 64 else (128 * nblocks,)
        t = s.sample_uniform_seeded_at(dtype, shape, seed, first_block).to_host()
        stream = np.zeros(16 * nblocks, dtype=np.uint8)
        sd = np.frombuffer(seed, dtype=np.uint8).copy()
        mo._aes_ctr(sd.ctypes.data_as(ctypes.c_void_p), first_block, nblocks, stream.ctypes.data_as(ctypes.c_void_p))
        words = stream.view(np.uint64).reshape(nblocks, 2)
        if bits == 128:      # element = (first u64 << 64) + second u64 (host/ops.rs:1998-2000): [lo, hi] = [second, first]
            want = words[:, ::-1]
            assert np.array_equal(t, want)
        elif bits == 64:
            assert np.array_equal(t, words.reshape(-1))
        else:
            assert np.array_equal(t, stream)
        if first_block == 0:
            assert np.array_equal(t, s.sample_uniform_seeded(dtype, shape, seed).to_host())
    finally:
        s.close()
