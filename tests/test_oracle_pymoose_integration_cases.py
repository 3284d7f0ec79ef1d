"""The reference's pymoose integration tests (pymoose/rust_integration_tests/*_test.py), at the protocol level on the
oracle (CPU): the same inputs, the same fixed-point types -- pymoose's `fixed(i, f)` always lowers to Fixed128
(pymoose/src/computation.rs:678-685), i.e. Ring128 -- and the same numpy answers at the same number of decimals.

  relu_test.py:36-57 (5 decimals)      sigmoid_test.py:55-73 (5)     exp_test.py:55-69 (5)
  log_test.py:36-56 (ln and log2, 5)   sqrt_test.py:36-49 (5)        softmax_test.py:54-108 (2)
  argmax_test.py:44-86 (exact)         reduce_max_test.py:41-62      maximum_test.py:102-139 (exact)
  add_n_test.py:20-60 (exact)          mirrored_ops_test.py:18-43
"""
import numpy as np
import pytest

from oracle import moose_oracle as mo

KEY = bytes(range(16))
BITS = 128


def shared(sess, x, frac, owner=0):
    return mo.rep_share(sess, mo.fixed_encode(BITS, np.asarray(x, dtype=np.float64), frac), owner)


def opened(x, frac, receiver=0):
    return mo.fixed_decode(mo.rep_reveal(x, receiver), frac)


def close(got, want, decimal):
    np.testing.assert_almost_equal(got, want, decimal=decimal)


@pytest.mark.parametrize("x", [[-1, -2, 0, 2, 3], [-1.25, -2.84, 0.001, 2.39, 3.12], [10.1, 8.3, 1.1], [-0.01, -0.03],
                               [0.02, -0.04, 1.08, -1.38, 3.65, -1.56],
                               [-0.71, 2.3, -0.74, 0.02, -0.04, 1.08, -0.64, 0.76, 0.97]])
def test_relu(x):
    sess = mo.Session("pymoose", KEY)
    close(opened(mo.rep_relu(sess, shared(sess, x, 27)), 27), np.maximum(np.array(x, dtype=np.float64), 0), 5)


@pytest.mark.parametrize("x", [[1, 0, 2, 3], [-1, 0, -2, -3.5], [-4.132, 0, -2, -3.5]])
def test_sigmoid_and_exp(x):
    sess = mo.Session("pymoose", KEY)
    v = np.array(x, dtype=np.float64)
    close(opened(mo.rep_sigmoid(sess, 8, 27, shared(sess, v, 27)), 27), 1.0 / (1.0 + np.exp(-v)), 5)
    close(opened(mo.rep_exp(sess, 8, 27, shared(sess, v, 27)), 27), np.exp(v), 5)


@pytest.mark.parametrize("x,fn,ref", [([1, 3, 2, 3], mo.rep_log, np.log), ([1.32, 10.42, 2.321, 3.5913], mo.rep_log, np.log),
                                      ([4.132, 1.932, 2, 4.5321], mo.rep_log, np.log),
                                      ([1, 2, 4, 8, 4.5, 10.5], mo.rep_log2, np.log2),
                                      ([1.0, 2.0, 3.0, 4.0], mo.rep_sqrt, np.sqrt), ([5.0, 6.0, 7.0, 8.0], mo.rep_sqrt, np.sqrt)])
def test_log_log2_sqrt(x, fn, ref):
    sess = mo.Session("pymoose", KEY)
    v = np.array(x, dtype=np.float64)
    close(opened(fn(sess, 8, 27, shared(sess, v, 27)), 27), ref(v), 5)


@pytest.mark.parametrize("x,axis,upmost", [
    ([[[1, 2, 3], [4, 5, 6]], [[7, 8, 9], [10, 11, 12]]], 0, 2),
    ([[[1, 10, 1], [4, 10, 3]], [[1, 5, 1], [3, 4, 3]]], 0, 2),
    ([[-1.38, 3.65, -1.56], [-1.38, 3.65, -1.8], [-0.64, 0.76, 0.97]], 1, 3),
    ([[-0.71, 2.3, -0.74], [-0.71, 2.3, -0.74], [0.02, -0.04, 1.08]], 1, 3),
    ([[-1.35, -0.63, -1.37], [-1.35, -0.63, -1.26], [-1.35, -1.34, -0.72]], 1, 3)])
def test_softmax(x, axis, upmost):
    sess = mo.Session("pymoose", KEY)
    v = np.array(x, dtype=np.float64)
    ex = np.exp(v - v.max(axis=axis, keepdims=True))
    close(opened(mo.rep_softmax(sess, axis, upmost, 8, 27, shared(sess, v, 27)), 27), ex / ex.sum(axis=axis, keepdims=True), 2)


@pytest.mark.parametrize("x,axis,upmost", [([[[1, 2, 3], [4, 5, 6]], [[7, 8, 9], [10, 11, 12]]], 0, 2),
                                           ([[[1, 10, 1], [4, 10, 3]], [[1, 5, 1], [3, 4, 3]]], 0, 2),
                                           ([-142.412, 0, 0, 0, 0], 0, 5)])
def test_argmax(x, axis, upmost):
    sess = mo.Session("pymoose", KEY)
    v = np.array(x, dtype=np.float64)
    got = mo.rep_reveal(mo.rep_argmax(sess, axis, upmost, shared(sess, v, 27))).to_ints()
    assert np.array_equal(np.asarray(got, dtype=np.int64), np.argmax(v, axis=axis))


@pytest.mark.parametrize("x", [[[[1, 2, 3], [4, 5, 6]], [[7, 8, 9], [10, 11, 12]]],
                               [[[1, 10, 1], [4, 100, 32]], [[123, 521, 132], [312, 421, 321]]]])
def test_reduce_max_over_the_last_axis(x):
    """index_axis(axis=2, i) on bob, maximum([x0, x1, x2]) on the replicated placement."""
    sess = mo.Session("pymoose", KEY)
    v = np.array(x, dtype=np.float64)
    enc = mo.fixed_encode(BITS, v, 27)
    parts = [mo.rep_share(sess, mo.index_axis(2, i, enc), 1) for i in range(3)]
    close(opened(mo.rep_maximum(sess, parts), 27, receiver=1), v.max(axis=2), 7)


def test_maximum_of_three_owners_and_add_n():
    sess = mo.Session("pymoose", KEY)
    x, y, z = np.array([1.0, 2.0, 9.0]), np.array([1.0, 2.0, 3.0]), np.array([4.0, 5.0, 6.0])
    parts = [shared(sess, x, 23, 0), shared(sess, y, 23, 1), shared(sess, z, 23, 2)]
    assert np.array_equal(opened(mo.rep_maximum(sess, parts), 23, receiver=1), np.maximum(x, np.maximum(y, z)))
    w = [shared(sess, [1.0, 2.0, 3.0], 23, 0), shared(sess, [1.0, 2.0, 3.0], 23, 0),
         shared(sess, [4.0, 5.0, 6.0], 23, 1), shared(sess, [7.0, 8.0, 9.0], 23, 2)]
    assert np.array_equal(opened(mo.rep_add_n(w), 23, receiver=1), np.array([13.0, 17.0, 21.0]))


def test_mirrored_constant_round_trip():
    """A mirrored fixed(8,27) constant read back on alice: the constant-to-ring conversion (kernels/constants.rs:59-130)
    and the decode."""
    v = np.array([1.5, 2.3, 3, 3], dtype=np.float64)
    ring = mo.Ring.from_ints(BITS, [mo.fixed_constant(BITS, float(t), 27) for t in v])
    close(mo.fixed_decode(ring, 27), v, 7)


def test_boolean_ops_less_greater_mux_with_bit_selector_and_concat():
    """boolean_ops_test.py:21-110: less / greater on the replicated placement, `mux(z_less, x, y)` with the replicated
    bit as selector (MuxOp::rep_bit_selector_kernel), and the concatenation of the bit tensor with itself."""
    sess = mo.Session("pymoose", KEY)
    x, y = np.array([1.5, 2.3, 3, 3]), np.array([-1.0, 4.0, 3, 2])
    xs, ys = shared(sess, x, 27, 1), shared(sess, y, 27, 1)
    less, greater = mo.rep_less(sess, xs, ys), mo.rep_greater(sess, xs, ys)
    assert mo.rep_reveal(less).to_array().tolist() == (x < y).astype(int).tolist()
    assert mo.rep_reveal(greater).to_array().tolist() == (x > y).astype(int).tolist()
    mux = mo.rep_mux_bit_selector(sess, BITS, less, xs, ys)
    close(opened(mux, 27), (x < y) * x + (1 - (x < y)) * y, 7)
    twice = mo.rep_concat(0, [less, less])
    assert mo.rep_reveal(twice).to_array().tolist() == np.concatenate([x < y, x < y]).astype(int).tolist()


@pytest.mark.parametrize("axis,expected", [
    (0, [[1.0, 2.0, 3.0], [1.0, 2.0, 3.0], [4.0, 5.0, 6.0], [7.0, 8.0, 9.0]]),
    (1, [[1.0, 2.0, 3.0, 1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0, 8.0, 9.0]])])
def test_concat_of_four_owners(axis, expected):
    """concat_test.py:19-72."""
    sess = mo.Session("pymoose", KEY)
    parts = [shared(sess, [[1.0, 2.0, 3.0]], 23, 0), shared(sess, [[1.0, 2.0, 3.0]], 23, 0),
             shared(sess, [[4.0, 5.0, 6.0]], 23, 1), shared(sess, [[7.0, 8.0, 9.0]], 23, 2)]
    assert np.array_equal(opened(mo.rep_concat(axis, parts), 23, receiver=1), np.array(expected))


def test_share_wise_shape_operators():
    """reshape_test.py:20-50 ([[1,2],[3,4]] -> [1,4]), slicing_test.py:120-146 (x[1:, 1:2]), squeeze_test.py and
    transpose_test.py on the replicated placement: share-wise, so the opened value is numpy's."""
    sess = mo.Session("pymoose", KEY)
    x = np.array([[1.0, 2.0], [3.0, 4.0]])
    xs = shared(sess, x, 27, 1)
    reshaped = mo.RepTensor([[v.reshape((1, 4)) for v in row] for row in xs.shares])
    assert np.array_equal(opened(reshaped, 27, 1), x.reshape(1, 4))
    assert np.array_equal(opened(mo.rep_transpose(xs), 27, 1), x.T)
    v = np.array([[1, 23.0, 321, 30.321, 321], [32.0, 321, 5, 3.0, 32.0]])
    vs = shared(sess, v, 27, 1)
    sliced = mo.RepTensor([[mo.slice_ring([(1, None, None), (1, 2, None)], t) for t in row] for row in vs.shares])
    close(opened(sliced, 27, 1), v[1:, 1:2], 7)
    one = shared(sess, np.array([[[1.0], [2.0]]]), 23, 0)
    assert opened(mo.rep_squeeze(None, one), 23).shape == (2,)
    assert opened(mo.rep_squeeze(2, one), 23).shape == (1, 2)
