"""mb_dot_noalloc: the ring GEMM variant for hosts that cannot let the library allocate device memory
(pre-partitioned HBM, latency-critical request paths).  It works out of one fixed array inside the
library instead of growable scratch buffers, so its results must not depend on what an earlier call left
in that array.  This file is named so that pytest collects it last: if it fails under `-x`, every other
test has already run.

Each case calls the new entry point twice in a row through the C ABI -- the second call with a smaller,
ragged K so that stale limb planes from the first call sit exactly where the second one needs zeros --
and compares every output bit with the regular mb_dot entry point on the same operands."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import RING64, RING128
from _util import ENTROPY_KEY, rand_ring

pytestmark = pytest.mark.gpu
DT = {64: RING64, 128: RING128}


def _dot_both_ways(sess, bits, M, K, N, seed):
    rng = np.random.default_rng(seed)
    a, b = rand_ring(rng, bits, (M, K)), rand_ring(rng, bits, (K, N))
    x, y = sess.from_host(DT[bits], a), sess.from_host(DT[bits], b)
    out = sess.alloc(DT[bits], (M, N))     # caller-owned output; the call under test allocates nothing
    sess.dot_noalloc(x, y, out)
    got = out.to_host()
    for mode in (2, 1):                    # existing entry point: tensor-pipe path, then the IMAD path
        sess.set_gemm_mode(mode)
        try:
            ref = sess.dot(x, y).to_host()
        finally:
            sess.set_gemm_mode(0)
        assert np.array_equal(got, ref), f"mb_dot_noalloc differs from mb_dot (gemm mode {mode})"


@pytest.mark.parametrize("bits", [64, 128])
def test_two_calls_in_a_row_match_mb_dot(bits):
    sess = mb.Session(0, "b200-noalloc", ENTROPY_KEY)
    try:
        _dot_both_ways(sess, bits, 256, 512, 256, seed=bits)       # fills the array up to its stated example size
        _dot_both_ways(sess, bits, 130, 200, 66, seed=bits + 1)    # ragged: K padded to 256, stale planes underneath
        _dot_both_ways(sess, bits, 256, 512, 256, seed=bits + 2)   # and back
    finally:
        sess.close()


def test_rejects_shapes_beyond_the_builtin_scratch():
    sess = mb.Session(0, "b200-noalloc", ENTROPY_KEY)
    try:
        M = K = N = 1024
        x = sess.fill(RING128, 1, (M, K))
        y = sess.fill(RING128, 1, (K, N))
        out = sess.alloc(RING128, (M, N))
        with pytest.raises(mb.MooseB200Error) as e:
            sess.dot_noalloc(x, y, out)
        assert e.value.code == -1 and "8388608" in str(e.value)
        with pytest.raises(mb.MooseB200Error):
            sess.dot_noalloc(sess.fill(RING64, 1, (64, 128)), sess.fill(RING64, 1, (128, 64)),
                             sess.alloc(RING64, (64, 64)))  # M below the tile size
    finally:
        sess.close()
