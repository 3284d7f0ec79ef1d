"""GPU parity tests of the shape / indexing rows of the HostPlacement kernel table (SURVEY 8a row a19 and the
rest of the 8b surface): through the C ABI, bit for bit against the oracle, plus the vectors of the reference's
own unit tests (embedded: the reference tree does not exist on the GPU box)."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import BIT, RING64, RING128
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY, SESSION_ID, rand_ring, same_bits, same_ring

pytestmark = pytest.mark.gpu

DT = {64: RING64, 128: RING128}


@pytest.fixture(scope="module")
def sess():
    s = mb.Session(0, SESSION_ID, ENTROPY_KEY)
    yield s
    s.close()


def ring(sess, bits, values):
    o = mo.Ring.from_ints(bits, values)
    return sess.from_host(DT[bits], o.data), o


def rnd(sess, rng, bits, shape):
    a = rand_ring(rng, bits, shape)
    return sess.from_host(DT[bits], a), mo.Ring(bits, a)


def bit_pair(sess, values):
    o = mo.Bits.from_array(np.asarray(values, dtype=np.uint8))
    return sess.from_bits(np.asarray(values, dtype=np.uint8)), o


def ints(t):
    return [int(v) for v in t.to_ints().reshape(-1)]


def test_reference_vectors(sess):
    # host/mod.rs:613-663 (test_tensor_slice, test_tensor_slice_neg_indicies)
    X, _ = ring(sess, 64, [[1, 2], [3, 4]])
    y = sess.slice([(1, None, None), (0, None, None)], X)
    assert y.shape == (1, 2) and ints(y) == [3, 4]
    y = sess.slice([(-1, None, 2), (-1, None, 2)], X)
    assert y.shape == (1, 1) and ints(y) == [4]
    with pytest.raises(mb.MooseB200Error):  # host/mod.rs:665-690: rank mismatch
        sess.slice([(-1, None, 2)] * 3, X)
    # host/mod.rs:693-714 (test_diag, test_bit_diag)
    assert ints(sess.diag(X)) == [1, 4]
    B, _ = bit_pair(sess, [[1, 1], [1, 0]])
    d = sess.diag(B)
    assert d.shape == (2,) and d.bits().tolist() == [1, 0]
    # host/mod.rs:717-761 (test_index, test_index_bit)
    X3, _ = ring(sess, 64, [[[1, 2], [3, 4]], [[4, 5], [6, 7]]])
    y = sess.index_axis(0, 1, X3)
    assert y.shape == (2, 2) and ints(y) == [4, 5, 6, 7]
    V, _ = ring(sess, 64, [0, 1, 2, 3])
    y = sess.expand_dims([0], sess.index_axis(0, 1, V))
    assert y.shape == (1,) and ints(y) == [1]
    B3, _ = bit_pair(sess, [[[0, 1], [0, 0]], [[1, 1], [0, 0]]])
    assert sess.index_axis(0, 1, B3).bits().tolist() == [[1, 1], [0, 0]]
    assert sess.index_axis(0, 0, B3).bits().tolist() == [[0, 1], [0, 0]]
    # host/mod.rs:764-787 (test_transpose, test_concatenate) on ring values
    assert ints(sess.transpose(X)) == [1, 3, 2, 4]
    C = sess.concat(0, [sess.expand_dims([0], X), sess.expand_dims([0], X)])
    assert C.shape == (2, 2, 2) and ints(C) == [1, 2, 3, 4, 1, 2, 3, 4]
    # host/mod.rs:1232-1288 (bit_transpose, bit_transpose2, bit_transpose3)
    xb = sess.bit_decompose(sess.from_host(RING64, np.array([6743216615002642708], dtype=np.uint64)))
    xt = sess.transpose(xb)
    expect = [0, 0, 1, 0, 1, 0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 1, 0, 1, 1, 1, 1, 1, 0, 1, 0, 0, 1, 0,
              0, 0, 1, 1, 0, 0, 0, 1, 0, 0, 1, 1, 1, 1, 1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 1, 0, 0, 1,
              1, 0, 1, 1, 1, 0, 1, 0]
    assert xt.shape == (1, 64) and xt.bits().reshape(-1).tolist() == expect
    B2, _ = bit_pair(sess, [[1, 1, 1], [0, 0, 0]])
    assert sess.transpose(B2).bits().tolist() == [[1, 0], [1, 0], [1, 0]]
    a3 = np.array([[[1, 1, 1], [0, 0, 0]], [[0, 1, 0], [1, 0, 0]]], dtype=np.uint8)
    B3, _ = bit_pair(sess, a3)
    assert sess.transpose(B3).bits().tolist() == a3.T.tolist()
    # host/fixedpoint.rs:120-145 (fixedpoint_mean_with_axis / _no_axis)
    E = sess.fixedpoint_encode(RING64, 16, sess.from_host(mb.FLOAT64, np.array([[1., 2.], [3., 4.]])))
    m = sess.fixedpoint_decode(32, sess.fixedpoint_mean(0, 2, 16, E)).to_host()
    assert m.tolist() == [2.0, 3.0]
    m = sess.fixedpoint_decode(32, sess.fixedpoint_mean(None, 2, 16, E)).to_host()
    assert float(m) == 2.5


@pytest.mark.parametrize("bits", [64, 128])
def test_expand_squeeze_transpose_parity(sess, bits):
    rng = np.random.default_rng(70 + bits)
    X, ox = rnd(sess, rng, bits, (3, 1, 5))
    for axes in ([0], [3], [1, 0], [0, 2, 4]):
        assert same_ring(sess.expand_dims(axes, X), mo.expand_dims(axes, ox))
    assert same_ring(sess.squeeze(None, X), mo.squeeze(None, ox))
    assert same_ring(sess.squeeze(1, X), mo.squeeze(1, ox))
    with pytest.raises(mb.MooseB200Error):
        sess.squeeze(0, X)
    for shape in [(7,), (37, 65), (1, 9), (5, 4, 3), (2, 3, 4, 5), (0, 4), ()]:
        X, ox = rnd(sess, rng, bits, shape)
        assert same_ring(sess.transpose(X), mo.transpose(ox))


def test_bit_shape_ops_parity(sess):
    rng = np.random.default_rng(77)
    for shape in [(5,), (33, 70), (3, 5, 7), (64, 130)]:
        a = rng.integers(0, 2, shape).astype(np.uint8)
        B, ob = bit_pair(sess, a)
        assert same_bits(sess.transpose(B), mo.transpose(ob))
        assert same_bits(sess.expand_dims([0], B), mo.expand_dims([0], ob))
        for axis in range(len(shape)):
            idx = shape[axis] // 2
            assert same_bits(sess.index_axis(axis, idx, B), mo.index_axis(axis, idx, ob))
    parts = [rng.integers(0, 2, (n, 13)).astype(np.uint8) for n in (3, 1, 40)]
    Bs, obs = zip(*(bit_pair(sess, p) for p in parts))
    assert same_bits(sess.concat(0, list(Bs)), mo.concat(0, list(obs)))
    with pytest.raises(mb.MooseB200Error):
        sess.transpose(sess.from_bits(np.zeros((2, 2, 2, 2), dtype=np.uint8)))
    with pytest.raises(mb.MooseB200Error):  # bitarray.rs:130-133
        sess.slice([(0, None, None)], sess.from_bits(np.zeros((4,), dtype=np.uint8)))
    with pytest.raises(mb.MooseB200Error) as e:  # bitarray.rs:124 todo!()
        sess.diag(sess.from_bits(np.zeros((3, 3), dtype=np.uint8)))
    assert e.value.code == -3


@pytest.mark.parametrize("bits", [64, 128])
def test_concat_index_slice_select_parity(sess, bits):
    rng = np.random.default_rng(90 + bits)
    X, ox = rnd(sess, rng, bits, (4, 6, 5))
    Y, oy = rnd(sess, rng, bits, (4, 2, 5))
    Z, oz = rnd(sess, rng, bits, (1, 6, 5))
    assert same_ring(sess.concat(1, [X, Y, X]), mo.concat(1, [ox, oy, ox]))
    assert same_ring(sess.concat(0, [Z, X]), mo.concat(0, [oz, ox]))
    W, ow = rnd(sess, rng, bits, (4, 6, 1))
    assert same_ring(sess.concat(2, [W, X]), mo.concat(2, [ow, ox]))  # the bias-trick concat of linear predictors
    with pytest.raises(mb.MooseB200Error):
        sess.concat(0, [X, Y])
    for axis in range(3):
        for idx in (0, ox.shape[axis] - 1):
            assert same_ring(sess.index_axis(axis, idx, X), mo.index_axis(axis, idx, ox))
    with pytest.raises(mb.MooseB200Error) as e:
        sess.index_axis(1, 6, X)
    assert e.value.code == -1
    for info in ([(0, None, None), (1, 5, 2), (0, None, None)], [(-3, -1, 1), (0, 6, -1), (4, None, None)],
                 [(1, 1, None), (0, None, 3), (0, None, -2)], [(0, 4, -3), (5, None, -1), (-5, None, 2)]):
        assert same_ring(sess.slice(info, X), mo.slice_ring(info, ox))
    with pytest.raises(mb.MooseB200Error):
        sess.slice([(0, 7, None), (0, None, None), (0, None, None)], X)
    for pick in ([1, 0, 1, 1, 0, 1], [0] * 6, [1] * 6):
        I, oi = bit_pair(sess, pick)
        assert same_ring(sess.select(1, I, X), mo.select(1, oi, ox))
    D, od = rnd(sess, rng, bits, (5, 7))
    assert same_ring(sess.diag(D), mo.diag(od))
    D3, od3 = rnd(sess, rng, bits, (3, 4, 5))
    assert same_ring(sess.diag(D3), mo.diag(od3))


@pytest.mark.parametrize("bits", [64, 128])
def test_div_planes_mean_argmax_parity(sess, bits):
    rng = np.random.default_rng(110 + bits)
    X, ox = rnd(sess, rng, bits, (6, 9))
    small = rng.integers(1, 1 << 40, (6, 9)).astype(np.uint64)
    if bits == 128:
        small = np.stack([small, rng.integers(0, 3, (6, 9)).astype(np.uint64)], axis=-1)
    Y, oy = sess.from_host(DT[bits], small), mo.Ring(bits, small)
    assert same_ring(sess.div(X, Y), mo.ring_div(ox, oy))
    V, ov = ring(sess, bits, [3, 5, 7, 11, 13, 17, 19, 23, 29])
    assert same_ring(sess.div(X, V), mo.ring_div(ox, ov))  # ndarray broadcasting
    with pytest.raises(mb.MooseB200Error) as e:
        sess.div(X, sess.fill(DT[bits], 0, (6, 9)))
    assert e.value.code == -2
    P, op = rnd(sess, rng, bits, (3, 5))
    assert same_ring(sess.bit_decompose_ring(P), mo.bit_decompose_ring(op))
    F = rng.uniform(-100, 100, (8, 5, 3))
    frac = 23 if bits == 64 else 40
    E, oe = sess.fixedpoint_encode(DT[bits], frac, sess.from_host(mb.FLOAT64, F)), mo.fixed_encode(bits, F, frac)
    for axis in (None, 0, 1, 2):
        assert same_ring(sess.fixedpoint_mean(axis, 2, frac, E), mo.fixedpoint_mean(axis, 2, frac, oe))
    for axis in (0, 1, 2):
        assert same_ring(sess.fixedpoint_argmax(axis, E), mo.fixedpoint_argmax(axis, oe))
    T, ot = ring(sess, bits, [[1, 5, 5], [-1, -3, -2]])
    assert ints(sess.fixedpoint_argmax(1, T)) == [1, 0]


def test_large_transpose_and_concat(sess):
    """Sizes past one tile / one grid pass; checked by properties (involution, row content)."""
    rng = np.random.default_rng(5)
    a = rand_ring(rng, 128, (1000, 515))
    X = sess.from_host(RING128, a)
    XT = sess.transpose(X)
    assert np.array_equal(XT.to_host(), np.ascontiguousarray(np.swapaxes(a, 0, 1)))
    assert np.array_equal(sess.transpose(XT).to_host(), a)
    ones = sess.fill(RING128, 1 << 40, (1000, 1))
    C = sess.concat(1, [ones, X]).to_host()
    assert C.shape == (1000, 516, 2) and np.array_equal(C[:, 1:], a) and np.all(C[:, 0, 0] == 1 << 40)
