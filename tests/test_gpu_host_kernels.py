"""GPU parity tests of the HostPlacement kernel table: every call goes through the C ABI
(ctypes) and is compared bit-for-bit with the oracle on the same seeded inputs.  Golden
vectors cited from /root/reference/moose/src are embedded (the reference tree does not
exist on the GPU box)."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import BIT, FLOAT64, RING64, RING128
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY, SESSION_ID, rand_ring, same_bits, same_ring

pytestmark = pytest.mark.gpu

DT = {64: RING64, 128: RING128}


@pytest.fixture(scope="module")
def sess():
    s = mb.Session(0, SESSION_ID, ENTROPY_KEY)
    yield s
    s.close()


def ints(t):
    return [int(v) for v in t.to_ints().reshape(-1)]


def test_ring_sample_golden(sess):
    """host/mod.rs:1050-1085 (ring_sample)."""
    seed = bytes(16)
    assert ints(sess.sample_uniform_seeded(RING64, (5,), seed)) == [
        4263935709876578662, 3326810793440857224, 17325099178452873543,
        15208531650305571673, 9619880027406922172]
    assert ints(sess.sample_uniform_seeded(RING128, (5,), seed)) == [
        78655730786844307471556614669614075016, 319591670596555766473793801091584867161,
        177455464885365520564027128957528354027, 72628979995024532377123578937486303732,
        299726520301515014350190124791858941972]
    assert ints(sess.sample_bits_seeded(RING64, (5,), seed)) == [0, 1, 1, 0, 0]
    assert ints(sess.sample_bits_seeded(RING128, (5,), seed)) == [0, 1, 1, 0, 0]
    # host/mod.rs:941-951 (bit_sample)
    assert sess.sample_uniform_seeded(BIT, (5,), seed).bits().tolist() == [0, 1, 1, 0, 0]


def test_sample_unsupported_max_value(sess):
    """host/ops.rs:1975-1978, 2025-2029: UnimplementedOperator."""
    with pytest.raises(mb.MooseB200Error) as e:
        sess.sample_uniform_seeded(RING64, (4,), bytes(16), max_value=7)
    assert e.value.code == -3
    with pytest.raises(mb.MooseB200Error) as e:
        sess.sample_uniform_seeded(BIT, (4,), bytes(16), max_value=1)
    assert e.value.code == -3


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("shape", [(), (1,), (7,), (3, 5), (257, 33), (4, 3, 2, 5), (0,), (100003,)])
def test_sample_seeded_stream_parity(sess, bits, shape):
    seed = bytes((b * 7 + 3) % 256 for b in range(16))
    assert same_ring(sess.sample_uniform_seeded(DT[bits], shape, seed), mo.sample_uniform_seeded(bits, shape, seed))


@pytest.mark.parametrize("shape", [(5,), (128,), (129,), (64, 37), (1000003,), (0,)])
def test_sample_bits_parity(sess, shape):
    seed = bytes(range(100, 116))
    assert same_bits(sess.sample_uniform_seeded(BIT, shape, seed), mo.sample_uniform_seeded_bits(shape, seed))
    if int(np.prod(shape)) < 5000:
        for bits in (64, 128):
            assert same_ring(sess.sample_bits_seeded(DT[bits], shape, seed), mo.sample_bits_seeded_ring(bits, shape, seed))


def test_derive_seed_and_session_id(sess):
    """execution/mod.rs:172-199 via the context's session id."""
    s2 = mb.Session(0, "foobar", ENTROPY_KEY)
    try:
        assert s2.session_id_raw() == mo.session_id_bytes("foobar")
        seed = s2.derive_seed(mo.sync_key_from([1, 2, 3]), bytes(16))
        assert list(seed) == [79, 203, 243, 208, 77, 199, 116, 216, 2, 206, 173, 36, 20, 204, 200, 146]
    finally:
        s2.close()


def test_entropy_source_contract():
    a = mb.Session(0, SESSION_ID, ENTROPY_KEY)
    try:
        o = mo.Session(SESSION_ID, ENTROPY_KEY)
        assert a.gen_key() == o.gen_key()
        assert a.sync_key_random() == o.sync_key_random()
        t = a.sample_uniform(RING64, (9,))
        assert same_ring(t, o.sample_uniform(mo.fill(64, 0, (9,))))
    finally:
        a.close()


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("shape", [(), (1,), (37,), (5, 7), (129, 65), (2, 3, 4, 5)])
def test_elementwise_parity(sess, bits, shape):
    rng = np.random.default_rng(hash((bits, shape)) % 2 ** 32)
    a, b = rand_ring(rng, bits, shape), rand_ring(rng, bits, shape)
    A, B = sess.from_host(DT[bits], a), sess.from_host(DT[bits], b)
    oa, ob = mo.Ring(bits, a), mo.Ring(bits, b)
    assert same_ring(sess.add(A, B), mo.add(oa, ob))
    assert same_ring(sess.sub(A, B), mo.sub(oa, ob))
    assert same_ring(sess.mul(A, B), mo.mul(oa, ob))
    assert same_ring(sess.neg(A), mo.neg(oa))
    assert same_ring(sess.and_(A, B), mo.ring_and(oa, ob))
    for amount in (0, 1, 13, 40, 63, 64, 65, 127, 128):
        assert same_ring(sess.shl(amount, A), mo.shl(amount, oa))
        assert same_ring(sess.shr(amount, A), mo.shr(amount, oa))


def test_host_mul_u128_wrap(sess):
    """host/mod.rs:1291-1299 (test_host_mul)."""
    x = sess.from_ints(RING128, [340282366920938463463374415046855271599])
    assert ints(sess.mul(x, x)) == [37011954726876357358499180449]


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("sx,sy", [((4, 5), (5,)), ((4, 5), (1, 5)), ((4, 1), (1, 5)), ((3, 4, 5), (4, 1)),
                                   ((1,), (6, 2)), ((), (3,))])
def test_broadcasting_parity(sess, bits, sx, sy):
    """ndarray co-broadcast in x.0 + y.0 (host/mod.rs:813-829)."""
    rng = np.random.default_rng(5)
    a, b = rand_ring(rng, bits, sx), rand_ring(rng, bits, sy)
    A, B = sess.from_host(DT[bits], a), sess.from_host(DT[bits], b)
    oa, ob = mo.Ring(bits, a), mo.Ring(bits, b)
    assert same_ring(sess.add(A, B), mo.add(oa, ob))
    assert same_ring(sess.sub(A, B), mo.sub(oa, ob))
    assert same_ring(sess.mul(A, B), mo.mul(oa, ob))
    assert same_ring(sess.broadcast(B, np.broadcast_shapes(sx, sy)), mo.broadcast(ob, np.broadcast_shapes(sx, sy)))


def test_broadcast_mismatch_is_kernel_error(sess):
    a = sess.fill(RING64, 1, (3, 4))
    b = sess.fill(RING64, 1, (5,))
    with pytest.raises(mb.MooseB200Error) as e:
        sess.add(a, b)
    assert e.value.code == -2


def test_ring_dot_kats(sess):
    """host/mod.rs:1011-1047."""
    x = sess.from_ints(RING64, [[1, 2], [3, 4]])
    assert ints(sess.dot(x, sess.from_ints(RING64, [1, 1]))) == [3, 7]
    z = sess.dot(x, sess.from_ints(RING64, [[1, 0], [0, 1]]))
    assert z.shape == (2, 2) and ints(z) == [1, 2, 3, 4]
    z = sess.dot(sess.from_ints(RING64, [1, 2]), sess.from_ints(RING64, [1, 1]))
    assert z.shape == () and ints(z) == [3]


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("M,K,N", [(1, 1, 1), (3, 5, 4), (64, 64, 64), (65, 17, 66), (130, 100, 7), (200, 513, 2),
                                   (1000, 33, 10), (70, 129, 129)])
def test_dot_parity_imad(sess, bits, M, K, N):
    rng = np.random.default_rng(M * 1000 + K * 10 + N)
    a, b = rand_ring(rng, bits, (M, K)), rand_ring(rng, bits, (K, N))
    sess.set_gemm_mode(1)
    try:
        z = sess.dot(sess.from_host(DT[bits], a), sess.from_host(DT[bits], b))
    finally:
        sess.set_gemm_mode(0)
    assert same_ring(z, mo.dot(mo.Ring(bits, a), mo.Ring(bits, b)))


@pytest.mark.parametrize("bits", [64, 128])
def test_dot_rank1_variants(sess, bits):
    rng = np.random.default_rng(11)
    a, v, w = rand_ring(rng, bits, (6, 9)), rand_ring(rng, bits, (9,)), rand_ring(rng, bits, (6,))
    A, V, W = (sess.from_host(DT[bits], t) for t in (a, v, w))
    oa, ov, ow = mo.Ring(bits, a), mo.Ring(bits, v), mo.Ring(bits, w)
    assert same_ring(sess.dot(A, V), mo.dot(oa, ov))
    assert same_ring(sess.dot(W, A), mo.dot(ow, oa))
    assert same_ring(sess.dot(V, V), mo.dot(ov, ov))
    with pytest.raises(mb.MooseB200Error):
        sess.dot(A, W)  # inner dimensions differ
    with pytest.raises(mb.MooseB200Error):
        sess.dot(sess.fill(DT[bits], 1, (2, 2, 2)), A)  # rank 3


def test_add_n_and_sum(sess):
    """host/mod.rs:1119-1148 (ring_add_n)."""
    x = sess.from_ints(RING64, [[1, 4], [9, 16], [25, 36]])
    y = sess.from_ints(RING64, [[1, 3], [6, 10], [15, 21]])
    z = sess.from_ints(RING64, [[1, 36], [1225, 41616], [1413721, 48024900]])
    assert ints(sess.add_n([x])) == [1, 4, 9, 16, 25, 36]
    assert ints(sess.add_n([x, y, z])) == [3, 43, 1240, 41642, 1413761, 48024957]
    with pytest.raises(mb.MooseB200Error):
        sess.add_n([])
    rng = np.random.default_rng(3)
    for bits in (64, 128):
        a = rand_ring(rng, bits, (5, 6, 7))
        A, oa = sess.from_host(DT[bits], a), mo.Ring(bits, a)
        for axis in (None, 0, 1, 2):
            assert same_ring(sess.sum(A, axis), mo.sum_axis(oa, axis))


BITS_6743 = [0, 0, 1, 0, 1, 0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 1, 0, 1, 1, 1, 1, 1, 0, 1, 0, 0, 1, 0,
             0, 0, 1, 1, 0, 0, 0, 1, 0, 0, 1, 1, 1, 1, 1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 1, 0, 0, 1,
             1, 0, 1, 1, 1, 0, 1, 0]


def test_bit_kats(sess):
    """host/mod.rs:1151-1225 (bit_extract, bit_decompose1/2, bit_shl_dim)."""
    x = sess.from_ints(RING64, [6743216615002642708])
    xb = sess.bit_decompose(x)
    assert xb.shape == (64, 1) and xb.bits().reshape(-1).tolist() == BITS_6743
    assert sess.shl_dim(6, 64, xb).bits().reshape(-1).tolist() == [0] * 6 + BITS_6743[:58]
    v = sess.fill(RING64, 7, (5,))
    assert sess.bit_extract(2, v).bits().tolist() == [1] * 5
    assert sess.bit_extract(3, v).bits().tolist() == [0] * 5
    x = sess.from_ints(RING64, [[[1, 2], [3, 4]], [[4, 5], [6, 7]]])
    xb = sess.bit_decompose(x)
    for i in range(64):
        assert np.array_equal(sess.index_axis0(i, xb).to_host(), sess.bit_extract(i, x).to_host())


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("shape", [(1,), (5,), (128,), (3, 43), (256, 3), (1000,)])
def test_bit_ops_parity(sess, bits, shape):
    rng = np.random.default_rng(17)
    a = rand_ring(rng, bits, shape)
    A, oa = sess.from_host(DT[bits], a), mo.Ring(bits, a)
    gb, ob = sess.bit_decompose(A), mo.bit_decompose(oa)
    assert same_bits(gb, ob)
    for idx in (0, 1, bits // 2, bits - 1):
        assert same_bits(sess.bit_extract(idx, A), mo.bit_extract(idx, oa))
        assert same_bits(sess.index_axis0(idx, gb), mo.index_axis0(idx, ob))
    for amount in (0, 1, 2, 16, bits // 2, bits):
        assert same_bits(sess.shl_dim(amount, bits, gb), mo.shl_dim(amount, bits, ob))
    plane, oplane = sess.index_axis0(3, gb), mo.index_axis0(3, ob)
    for idx in (0, 5, bits - 1):
        assert same_ring(sess.ring_inject(DT[bits], idx, plane), mo.ring_inject(bits, idx, oplane))
    g2, o2 = sess.shl_dim(1, bits, gb), mo.shl_dim(1, bits, ob)
    assert same_bits(sess.xor(gb, g2), mo.bit_op("xor", ob, o2))
    assert same_bits(sess.and_(gb, g2), mo.bit_op("and", ob, o2))
    assert same_bits(sess.or_(gb, g2), mo.bit_op("or", ob, o2))
    assert same_bits(sess.neg(gb), mo.neg(ob))
    assert same_bits(sess.add(gb, g2), mo.add(ob, o2))  # Z2: add = xor
    assert same_bits(sess.mul(gb, g2), mo.mul(ob, o2))  # Z2: mul = and


def test_fill(sess):
    """host/mod.rs ring_fill / bit_fill."""
    assert ints(sess.fill(RING64, 7, (3,))) == [7, 7, 7]
    assert ints(sess.fill(RING128, (1 << 100) + 5, (2, 2))) == [(1 << 100) + 5] * 4
    assert sess.fill(BIT, 1, (2,)).bits().tolist() == [1, 1]
    assert np.array_equal(sess.fill(BIT, 1, (13,)).to_host(), mo.fill_bits(1, (13,)).packed)
    with pytest.raises(mb.MooseB200Error):
        sess.fill(BIT, 2, (2,))


@pytest.mark.parametrize("bits,frac", [(64, 23), (128, 40), (128, 27)])
def test_fixedpoint_encode_decode(sess, bits, frac):
    """host/ops.rs:1362-1418, host/fixedpoint.rs:10-38."""
    rng = np.random.default_rng(23)
    v = np.concatenate([rng.uniform(-1000, 1000, 500), [0.0, -0.0, 1.5, -1.5, 1e-30, -1e-30, 2.0 ** 30, -2.0 ** 30,
                                                          1e300, -1e300, np.inf, -np.inf, np.nan]])
    e = sess.fixedpoint_encode(DT[bits], frac, sess.from_host(FLOAT64, v))
    oe = mo.fixed_encode(bits, v, frac)
    assert same_ring(e, oe)
    d = sess.fixedpoint_decode(frac, e).to_host()
    assert np.array_equal(d, mo.fixed_decode(oe, frac))
    # round-to-nearest-even of wide integers (`i128 as f64`)
    w = rand_ring(rng, bits, (4096,))
    assert np.array_equal(sess.fixedpoint_decode(frac, sess.from_host(DT[bits], w)).to_host(),
                          mo.fixed_decode(mo.Ring(bits, w), frac))


RING_FIXEDPOINT = {(64, 16): [65536, 18446744073709420544, 196608, 18446744073709289472],
                   (128, 80): [1208925819614629174706176, 340282366920936045611735378173418799104,
                               3626777458843887524118528, 340282366920933627760096148915069386752]}


@pytest.mark.parametrize("bits,frac", list(RING_FIXEDPOINT))
def test_ring_fixedpoint_reference_vector(sess, bits, frac):
    """fixedpoint/ops.rs:2575-2627 (ring_fixedpoint): [1, -2, 3, -4] at 2^16 on Ring64 and at 2^80 on Ring128."""
    v = np.array([1.0, -2.0, 3.0, -4.0])
    e = sess.fixedpoint_encode(DT[bits], frac, sess.from_host(FLOAT64, v))
    assert ints(e) == RING_FIXEDPOINT[(bits, frac)]
    assert np.array_equal(sess.fixedpoint_decode(frac, e).to_host(), v)


@pytest.mark.parametrize("mode", [2, 3])
@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("M,K,N", [(128, 128, 64), (128, 256, 256), (256, 128, 128), (300, 200, 130), (129, 1000, 514),
                                   (512, 1024, 512), (700, 384, 258)])
def test_dot_parity_tensor_pipe(sess, bits, M, K, N, mode):
    """tcgen05 int8 limb GEMM against the oracle: full-range random ring elements, ragged M / K / N (TMA
    zero fill, padded K, predicated epilogue).  Mode 2 uses CTA pairs (cta_group::2, 256-row tiles) where
    the shape allows, mode 3 forces single-CTA tiles."""
    rng = np.random.default_rng(M * 7 + K * 3 + N)
    a, b = rand_ring(rng, bits, (M, K)), rand_ring(rng, bits, (K, N))
    sess.set_gemm_mode(mode)
    try:
        z = sess.dot(sess.from_host(DT[bits], a), sess.from_host(DT[bits], b))
    finally:
        sess.set_gemm_mode(0)
    assert same_ring(z, mo.dot(mo.Ring(bits, a), mo.Ring(bits, b)))


def test_dot_tensor_pipe_extreme_values(sess):
    """All-ones limbs (0xff..ff) maximise every u8 x u8 product: checks the accumulator-exactness
    planning (pass splitting) of the limb GEMM, K = 2048 > 66051 / 34."""
    M, K, N = 128, 2048, 64
    for bits in (64, 128):
        a = np.full((M, K) + ((2,) if bits == 128 else ()), 0xFFFFFFFFFFFFFFFF, dtype=np.uint64)
        b = np.full((K, N) + ((2,) if bits == 128 else ()), 0xFFFFFFFFFFFFFFFF, dtype=np.uint64)
        sess.set_gemm_mode(2)
        try:
            z = sess.dot(sess.from_host(DT[bits], a), sess.from_host(DT[bits], b))
        finally:
            sess.set_gemm_mode(0)
        assert same_ring(z, mo.dot(mo.Ring(bits, a), mo.Ring(bits, b)))


@pytest.mark.parametrize("bits", [64, 128])
def test_plaintext_helpers_parity(sess, bits):
    """host/ops.rs:2051-2087,2125-2161 (less/greater), :2227-2252 (mux), :2482-2506 (sign), :2277-2288 (cast),
    :1432-1452 (maximum, unsigned order)."""
    rng = np.random.default_rng(bits + 5)
    shape = (7, 19)
    a, b, c = (rand_ring(rng, bits, shape) for _ in range(3))
    small = (rng.integers(-5, 5, shape).astype(np.int64)).astype(np.uint64)
    if bits == 128:
        small = np.stack([small, np.where(small.astype(np.int64) < 0, np.uint64(0xFFFFFFFFFFFFFFFF), np.uint64(0))], axis=-1)
    A, B, C, S = (sess.from_host(DT[bits], t) for t in (a, b, c, small))
    oa, ob, oc, osm = (mo.Ring(bits, t) for t in (a, b, c, small))
    for (X, Y, ox, oy) in ((A, B, oa, ob), (S, sess.fill(DT[bits], 0, shape), osm, mo.fill(bits, 0, shape)), (A, A, oa, oa)):
        assert same_bits(sess.less(X, Y), mo.host_less(ox, oy))
        assert same_bits(sess.greater(X, Y), mo.host_greater(ox, oy))
    sel, osel = sess.less(A, B), mo.host_less(oa, ob)
    assert same_ring(sess.host_mux(sel, B, C), mo.host_mux(osel, ob, oc))
    assert same_ring(sess.sign(S), mo.host_sign(osm)) and same_ring(sess.sign(A), mo.host_sign(oa))
    assert same_ring(sess.host_maximum([A, B, C]), mo.host_maximum([oa, ob, oc]))
    with pytest.raises(mb.MooseB200Error):
        sess.host_maximum([])
    if bits == 128:
        assert same_ring(sess.cast_ring128_to_ring64(A), mo.cast_ring128_to_ring64(oa))
    # broadcasting compare: [7,19] vs [19]
    v = rand_ring(rng, bits, (19,))
    assert same_bits(sess.less(A, sess.from_host(DT[bits], v)), mo.host_less(oa, mo.Ring(bits, v)))


def test_async_transfers_across_contexts():
    """mb_tensor_from_host_async / mb_tensor_to_host_async / mb_ctx_wait: loader -> compute -> drainer contexts on one
    device give the same bytes as the synchronous path."""
    loader, compute, drainer = (mb.Session(0, SESSION_ID, ENTROPY_KEY) for _ in range(3))
    try:
        rng = np.random.default_rng(21)
        n = 1 << 20
        ha, hb, hout = (loader.pinned_empty((n, 2), np.uint64) for _ in range(3))
        ha[...] = rand_ring(rng, 128, (n,))
        hb[...] = rand_ring(rng, 128, (n,))
        for _ in range(3):
            a, b = loader.from_host_async(RING128, ha), loader.from_host_async(RING128, hb)
            compute.wait(loader)
            c = compute.mul(compute.add(a, b), b)
            c.sess = drainer
            drainer.wait(compute)
            drainer.to_host_async(c, hout)
            loader.wait(compute)
            del a, b
            drainer.sync()
            assert np.array_equal(hout, mo.mul(mo.add(mo.Ring(128, ha), mo.Ring(128, hb)), mo.Ring(128, hb)).data)
            del c
        with pytest.raises(mb.MooseB200Error):
            drainer.to_host_async(compute.fill(RING64, 1, (4,)), hout)
    finally:
        for s in (loader, compute, drainer):
            s.close()


def test_tensor_shapes_that_wrap_the_element_count_are_rejected(sess):
    """Shapes can come off the wire (a peer's header): [2^32, 2^32] must not wrap to 0 elements."""
    for shape in [(1 << 32, 1 << 32), (1 << 62, 4), (1 << 41,)]:
        with pytest.raises(mb.MooseB200Error):
            sess.alloc(RING64, shape)
    assert sess.alloc(RING64, (0, 5)).numel == 0
