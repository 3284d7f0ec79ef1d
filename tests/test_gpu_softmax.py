"""Softmax, argmax, abs and the share-wise shape operators of the replicated placement through
the C ABI: share-level parity with the oracle, and the vectors / tolerances of the reference's own tests
(moose/src/replicated/softmax.rs:148-300: abs error < 0.01 resp. 0.015; moose/src/replicated/argmax.rs:123-200)."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import RING64, RING128
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY, SESSION_ID, rand_ring, same_rep

pytestmark = pytest.mark.gpu
DT = {64: RING64, 128: RING128}


def pair():
    return mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)


def shared(g, o, bits, frac, vals, owner=0):
    v = np.asarray(vals, dtype=np.float64)
    gx = g.share(g.fixedpoint_encode(DT[bits], frac, g.from_host(mb.FLOAT64, v)), owner)
    ox = mo.rep_share(o, mo.fixed_encode(bits, v, frac), owner)
    return gx, ox


def shared_ints(g, o, bits, vals):
    r = mo.Ring.from_ints(bits, vals)
    return g.share(g.from_host(DT[bits], r.data), 0), mo.rep_share(o, r, 0)


def np_softmax(x, axis):
    e = np.exp(x - x.max(axis=axis, keepdims=True))
    return e / e.sum(axis=axis, keepdims=True)


SOFTMAX_CASES = [
    # (bits, axis, upmost, i, f, tolerance, x)  -- softmax.rs:189-193 and the tests below them
    (64, 0, 9, 8, 10, 0.01, [1, 2.5, -3.0, 4.0, 2.0, -2.0, -2.0, -3.0, 3.0]),
    (128, 0, 9, 8, 27, 0.01, [1, 2.5, -3.0, 4.0, 2.0, -2.0, -2.0, -3.0, 3.0]),
    (128, 1, 4, 24, 40, 0.015, [[0.90102809, 0.65720883, -0.02816407, 0.0535739], [0.61216721, 0.20281131, 1.7734221, -0.69106256],
                                [-0.08150293, -1.50330937, -0.99238243, -2.65759917]]),
    (128, 1, 3, 24, 40, 0.015, [[-1.35, -0.63, -1.37], [-1.35, -0.63, -1.26], [-1.35, -1.34, -0.72]]),
]


@pytest.mark.parametrize("bits,axis,upmost,i_p,f_p,tol,x", SOFTMAX_CASES)
def test_softmax_reference_vectors(bits, axis, upmost, i_p, f_p, tol, x):
    g, o = pair()
    x = np.array(x, dtype=np.float64)
    gx, ox = shared(g, o, bits, f_p, x)
    gz, oz = g.softmax(axis, upmost, i_p, f_p, gx), mo.rep_softmax(o, axis, upmost, i_p, f_p, ox)
    assert same_rep(gz, oz)
    got = g.fixedpoint_decode(f_p, g.reveal(gz)).to_host()
    assert got.shape == x.shape and np.abs(got - np_softmax(x, axis)).max() < tol
    g.close()


def test_softmax_batch_rows():
    """A [64, 10] batch of logits on Ring128 fixed(24,40) (MLPClassifier's output layer): parity and row sums."""
    g, o = pair()
    x = np.random.default_rng(3).normal(0, 2.0, (64, 10))
    gx, ox = shared(g, o, 128, 40, x)
    gz, oz = g.softmax(1, 10, 24, 40, gx), mo.rep_softmax(o, 1, 10, 24, 40, ox)
    assert same_rep(gz, oz)
    got = g.fixedpoint_decode(40, g.reveal(gz)).to_host()
    assert np.abs(got - np_softmax(x, 1)).max() < 0.02 and np.abs(got.sum(axis=1) - 1).max() < 0.05
    with pytest.raises(mb.MooseB200Error):
        g.softmax(2, 10, 24, 40, gx)
    g.close()


ARGMAX_CASES = [
    # argmax.rs:123-200
    (64, 0, 9, [1, 2, -3, 4, 2, 2, 2, 3, 105], 8),
    (64, 1, 4, [[1231, -323, -3, 12321], [93213, 12312321, -321, -3249], [3921, 4012, 3221, -321]], [3, 1, 1]),
    (64, 1, 3, [[-3233, 329423, 54843], [3994, 123, -31326], [-324, 55321, 23441]], [1, 0, 1]),
    (128, 0, 9, [1, 2, -3, 10000, 10000, 10000, 10000, 10000, 10000], 3),
    (128, 1, 4, [[1231, -323, -3, 12321], [93213, 12312321, -321, -3249], [3921, 4012, 3221, -321]], [3, 1, 1]),
    (128, 1, 3, [[-3233, 329423, 54843], [3994, 123, -31326], [-324, 55321, 23441]], [1, 0, 1]),
]


@pytest.mark.parametrize("bits,axis,upmost,x,expect", ARGMAX_CASES)
def test_argmax_reference_vectors(bits, axis, upmost, x, expect):
    g, o = pair()
    gx, ox = shared_ints(g, o, bits, x)
    gz, oz = g.argmax(axis, upmost, gx), mo.rep_argmax(o, axis, upmost, ox)
    assert same_rep(gz, oz)
    opened = g.reveal(gz)
    assert opened.dtype == RING64 and opened.to_ints().tolist() == expect
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
def test_abs_and_sharewise_ops(bits):
    g, o = pair()
    rng = np.random.default_rng(bits)
    vals = rng.integers(-1000, 1000, (4, 6)).tolist()
    gx, ox = shared_ints(g, o, bits, vals)
    ga, oa = g.rep_abs(gx), mo.rep_abs(o, ox)
    assert same_rep(ga, oa) and g.reveal(ga).to_ints().tolist() == np.abs(np.array(vals)).tolist()
    a = rand_ring(rng, bits, (3, 4, 5))
    gy, oy = g.share(g.from_host(DT[bits], a), 1), mo.rep_share(o, mo.Ring(bits, a), 1)
    assert same_rep(g.rep_index_axis(1, 2, gy), mo.rep_index_axis(1, 2, oy))
    assert same_rep(g.rep_expand_dims([0, 2], gy), mo.rep_expand_dims([0, 2], oy))
    assert same_rep(g.rep_squeeze(0, g.rep_expand_dims(0, gy)), oy)
    assert same_rep(g.rep_transpose(gy), mo.rep_transpose(oy))
    assert same_rep(g.rep_sum(gy, 1), mo.rep_sum(1, oy)) and same_rep(g.rep_sum(gy), mo.rep_sum(None, oy))
    assert same_rep(g.rep_concat(2, [gy, gy]), mo.rep_concat(2, [oy, oy]))
    assert same_rep(g.rep_add_n([gy, gy, gy]), mo.rep_add_n([oy, oy, oy]))
    assert same_rep(g.rep_reshape(gy, (12, 5)), mo._sharewise(lambda v: v.reshape((12, 5)), oy))
    assert same_rep(g.rep_slice([(0, None, 2), (1, 3, None), (0, None, -1)], gy),
                    mo._sharewise(lambda v: mo.slice_ring([(0, None, 2), (1, 3, None), (0, None, -1)], v), oy))
    assert same_rep(g.rep_diag(gy), mo._sharewise(mo.diag, oy))
    assert same_rep(g.rep_fixedpoint_mean(1, 2, 10, gy), mo.rep_fixedpoint_mean(1, 2, 10, oy))
    row = rand_ring(rng, bits, (1, 4, 1))
    gr, orr = g.share(g.from_host(DT[bits], row), 0), mo.rep_share(o, mo.Ring(bits, row), 0)
    assert same_rep(g.rep_broadcast(gr, (3, 4, 5)), mo.rep_broadcast(orr, (3, 4, 5)))
    if bits == 128:
        assert same_rep(g.rep_cast_ring128_to_ring64(gy), mo.rep_cast_64(oy))
    g.close()


def test_equal_reference_vector():
    """compare.rs:183-203 (test_equal)."""
    g, o = pair()
    gx, ox = shared_ints(g, o, 64, [1024, 5, 4])
    ry = mo.Ring.from_ints(64, [1024, 4, 5])
    gy, oy = g.share(g.from_host(RING64, ry.data), 1), mo.rep_share(o, ry, 1)
    ge, oe = g.rep_equal(gx, gy), mo.rep_equal(o, ox, oy)
    assert same_rep(ge, oe) and g.reveal(ge).bits().tolist() == [1, 0, 0]
    g.close()


LOG_CASES = [
    # (bits, i, f, fn, tolerance, x)   -- log.rs:258-298 and :330-432
    (64, 8, 20, "log2", 0.01, [1.0, 2.0, 4.0, 8.0, 4.5, 10.5]),
    (128, 10, 30, "log2", 0.01, [[1.0, 2.0, 4.0], [8.0, 4.5, 10.5]]),
    (64, 8, 20, "log", 0.01, [1.0, 2.5, 3.0, 4.0, 5.0]),
    (128, 10, 30, "log", 0.001, [[1.0, 2.5, 3.0, 4.0, 5.0]]),
    (128, 10, 30, "sqrt", 0.001, [0.25, 1.0, 2.0, 9.0, 100.0]),
    (64, 8, 20, "sqrt", 0.02, [0.25, 1.0, 2.0, 9.0]),
]


@pytest.mark.parametrize("bits,i_p,f_p,fn,tol,x", LOG_CASES)
def test_log_sqrt(bits, i_p, f_p, fn, tol, x):
    g, o = pair()
    x = np.array(x, dtype=np.float64)
    gx, ox = shared(g, o, bits, f_p, x)
    gz = getattr(g, "rep_" + fn)(i_p, f_p, gx)
    oz = getattr(mo, "rep_" + fn)(o, i_p, f_p, ox)
    assert same_rep(gz, oz)
    got = g.fixedpoint_decode(f_p, g.reveal(gz)).to_host()
    assert np.abs(got - getattr(np, fn)(x)).max() < tol
    g.close()


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("widths,n_out", [((1, 37), 2), ((5, 1, 64), 10), ((33,), 3), ((1, 40), 48)])
def test_dot_concat_equals_concat_then_dot(bits, widths, n_out):
    """mb_rep_dot_concat (fused skinny kernel over virtual parts; N = 48 takes the concat + dot fallback) against the
    oracle's ConcatOp + DotOp: same shares."""
    g, o = pair()
    rng = np.random.default_rng(sum(widths) + bits)
    rows = 130
    gx, ox = [], []
    for w, owner in zip(widths, (1, 0, 2)):
        a = rand_ring(rng, bits, (rows, w))
        gx.append(g.share(g.from_host(DT[bits], a), owner))
        ox.append(mo.rep_share(o, mo.Ring(bits, a), owner))
    b = rand_ring(rng, bits, (sum(widths), n_out))
    gy, oy = g.share(g.from_host(DT[bits], b), 0), mo.rep_share(o, mo.Ring(bits, b), 0)
    gz = g.rep_dot_concat(gx, gy)
    oz = mo.rep_dot(o, mo.rep_concat(1, ox), oy)
    assert same_rep(gz, oz)
    # the plain replicated dot takes the same fused kernel for skinny right operands
    assert same_rep(g.rep_dot(gx[0], g.share(g.from_host(DT[bits], b[: widths[0]]), 1)),
                    mo.rep_dot(o, ox[0], mo.rep_share(o, mo.Ring(bits, b[: widths[0]]), 1)))
    g.close()
