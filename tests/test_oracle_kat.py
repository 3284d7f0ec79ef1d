"""Pin the oracle against every golden vector the reference's own tests hold
for the hot path.  Paths cite /root/reference/moose/src."""
import numpy as np
import pytest

from oracle import moose_oracle as mo


def ints(t):
    return [int(v) for v in t.to_ints().reshape(-1)]


def test_aes128_fips197_vectors():
    # FIPS-197 appendix B and C.1
    assert mo.aes_block(bytes.fromhex("2b7e151628aed2a6abf7158809cf4f3c"),
                        bytes.fromhex("3243f6a8885a308d313198a2e0370734")).hex() == "3925841d02dc09fbdc118597196a0b32"
    assert mo.aes_block(bytes(range(16)), bytes.fromhex("00112233445566778899aabbccddeeff")).hex() \
        == "69c4e0d86a7b0430d8cdb78070b4c55a"


@pytest.mark.parametrize("portable", [0, 1])
def test_ring_sample_golden(portable):
    """host/mod.rs:1050-1079 (ring_sample): seed 0^16, shape [5]."""
    mo._force_portable(portable)
    try:
        seed = bytes(16)
        assert ints(mo.sample_uniform_seeded(64, (5,), seed)) == [
            4263935709876578662, 3326810793440857224, 17325099178452873543,
            15208531650305571673, 9619880027406922172]
        assert ints(mo.sample_uniform_seeded(128, (5,), seed)) == [
            78655730786844307471556614669614075016, 319591670596555766473793801091584867161,
            177455464885365520564027128957528354027, 72628979995024532377123578937486303732,
            299726520301515014350190124791858941972]
    finally:
        mo._force_portable(0)


@pytest.mark.parametrize("policy", [mo.BIT_POLICY_STREAM_LSB0, mo.BIT_POLICY_U16_PER_BIT])
def test_bit_sample_golden(policy):
    """host/mod.rs:941-951 (bit_sample) and :1081-1085 (sample_bits_seeded)."""
    seed = bytes(16)
    assert mo.sample_uniform_seeded_bits((5,), seed, policy).to_array().tolist() == [0, 1, 1, 0, 0]
    assert ints(mo.sample_bits_seeded_ring(64, (5,), seed, policy)) == [0, 1, 1, 0, 0]
    assert ints(mo.sample_bits_seeded_ring(128, (5,), seed, policy)) == [0, 1, 1, 0, 0]


def test_derive_seed_golden():
    """execution/mod.rs:172-199 (test_constants_derive_seed): key 0^16, sync_key [1,2,3],
    session id "foobar" (TestSyncExecutor default)."""
    sid = mo.session_id_bytes("foobar")
    seed = mo.derive_seed(sid, mo.sync_key_from([1, 2, 3]), bytes(16))
    assert list(seed) == [79, 203, 243, 208, 77, 199, 116, 216, 2, 206, 173, 36, 20, 204, 200, 146]


def test_host_mul_u128_wrap():
    """host/mod.rs:1291-1299 (test_host_mul)."""
    x = mo.Ring.from_ints(128, [340282366920938463463374415046855271599])
    assert ints(mo.mul(x, x)) == [37011954726876357358499180449]


def test_ring_dot_kats():
    """host/mod.rs:1011-1047."""
    x = mo.Ring.from_ints(64, [[1, 2], [3, 4]])
    assert ints(mo.dot(x, mo.Ring.from_ints(64, [1, 1]))) == [3, 7]
    z = mo.dot(x, mo.Ring.from_ints(64, [[1, 0], [0, 1]]))
    assert z.shape == (2, 2) and ints(z) == [1, 2, 3, 4]
    z = mo.dot(mo.Ring.from_ints(64, [1, 2]), mo.Ring.from_ints(64, [1, 1]))
    assert z.shape == () and ints(z) == [3]


def test_ring_add_n():
    """host/mod.rs:1119-1148."""
    x = mo.Ring.from_ints(64, [[1, 4], [9, 16], [25, 36]])
    y = mo.Ring.from_ints(64, [[1, 3], [6, 10], [15, 21]])
    z = mo.Ring.from_ints(64, [[1, 36], [1225, 41616], [1413721, 48024900]])
    assert ints(mo.add_n([x])) == [1, 4, 9, 16, 25, 36]
    assert ints(mo.add_n([x, y, z])) == [3, 43, 1240, 41642, 1413761, 48024957]
    w = mo.Ring.from_ints(128, [[6, 3, 10], [5, 16, 8], [4, 2, 1]])
    x = mo.Ring.from_ints(128, [[40, 20, 10], [5, 16, 8], [4, 2, 1]])
    y = mo.Ring.from_ints(128, [[42, 21, 64], [32, 16, 8], [4, 2, 1]])
    z = mo.Ring.from_ints(128, [[256, 128, 64], [32, 16, 8], [4, 2, 1]])
    assert ints(mo.add_n([w, x, y, z])) == [344, 172, 148, 74, 64, 32, 16, 8, 4]


BITS_6743 = [0, 0, 1, 0, 1, 0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 1, 0, 1, 1, 1, 1, 1, 0, 1, 0, 0, 1, 0,
             0, 0, 1, 1, 0, 0, 0, 1, 0, 0, 1, 1, 1, 1, 1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 1, 0, 0, 1,
             1, 0, 1, 1, 1, 0, 1, 0]


def test_bit_decompose_and_extract():
    """host/mod.rs:1151-1204 (bit_extract, bit_decompose1, bit_decompose2)."""
    x = mo.Ring.from_ints(64, [6743216615002642708])
    b = mo.bit_decompose(x)
    assert b.shape == (64, 1) and b.to_array().reshape(-1).tolist() == BITS_6743
    x = mo.Ring.from_ints(64, [[[1, 2], [3, 4]], [[4, 5], [6, 7]]])
    xb = mo.bit_decompose(x)
    for i in range(64):
        assert mo.index_axis0(i, xb) == mo.bit_extract(i, x)
    v = mo.fill(64, 7, (5,))
    assert mo.bit_extract(2, v).to_array().tolist() == [1] * 5
    assert mo.bit_extract(3, v).to_array().tolist() == [0] * 5


def test_bit_shl_dim():
    """host/mod.rs:1206-1225."""
    x = mo.Ring.from_ints(64, [6743216615002642708])
    out = mo.shl_dim(6, 64, mo.bit_decompose(x))
    assert out.to_array().reshape(-1).tolist() == [0] * 6 + BITS_6743[:58]


def test_shift_masks_amount_like_rust_wrapping():
    """Wrapping<T> << usize masks the amount (host/ops.rs:1864,1879)."""
    x = mo.Ring.from_ints(64, [5])
    assert ints(mo.shl(64, x)) == [5] and ints(mo.shr(64, x)) == [5]
    x = mo.Ring.from_ints(128, [5])
    assert ints(mo.shl(128, x)) == [5] and ints(mo.shl(127, x)) == [1 << 127]


def test_fixed_encode_decode():
    """host/ops.rs:1377-1417, host/fixedpoint.rs:10-38: truncation toward zero."""
    v = np.array([1.5, -1.5, 0.3, -0.3])
    e = mo.fixed_encode(128, v, 40)
    assert ints(e) == [int(t * 2 ** 40) % (1 << 128) for t in v]
    assert np.allclose(mo.fixed_decode(e, 40), v, atol=2 ** -40)
    e = mo.fixed_encode(64, v, 23)
    assert ints(e) == [int(t * 2 ** 23) % (1 << 64) for t in v]
    # fixedpoint/ops.rs:2575-2627 (ring_fixedpoint)
    v = np.array([1.0, -2.0, 3.0, -4.0])
    assert ints(mo.fixed_encode(64, v, 16)) == [65536, 18446744073709420544, 196608, 18446744073709289472]
    e = mo.fixed_encode(128, v, 80)
    assert ints(e) == [1208925819614629174706176, 340282366920936045611735378173418799104,
                       3626777458843887524118528, 340282366920933627760096148915069386752]
    assert np.array_equal(mo.fixed_decode(e, 80), v) and np.array_equal(mo.fixed_decode(mo.fixed_encode(64, v, 16), 16), v)


def test_elementwise_against_python_ints():
    rng = np.random.default_rng(0)
    for bits in (64, 128):
        mod = 1 << bits
        a = [int.from_bytes(rng.bytes(bits // 8), "little") for _ in range(37)]
        b = [int.from_bytes(rng.bytes(bits // 8), "little") for _ in range(37)]
        A, B = mo.Ring.from_ints(bits, a), mo.Ring.from_ints(bits, b)
        assert ints(mo.add(A, B)) == [(x + y) % mod for x, y in zip(a, b)]
        assert ints(mo.sub(A, B)) == [(x - y) % mod for x, y in zip(a, b)]
        assert ints(mo.mul(A, B)) == [(x * y) % mod for x, y in zip(a, b)]
        assert ints(mo.neg(A)) == [(-x) % mod for x in a]
        assert ints(mo.shl(13, A)) == [(x << 13) % mod for x in a]
        assert ints(mo.shr(13, A)) == [x >> 13 for x in a]
        M, K, N = 3, 5, 4
        a = [int.from_bytes(rng.bytes(bits // 8), "little") for _ in range(M * K)]
        b = [int.from_bytes(rng.bytes(bits // 8), "little") for _ in range(K * N)]
        C = mo.dot(mo.Ring.from_ints(bits, a, (M, K)), mo.Ring.from_ints(bits, b, (K, N)))
        exp = [sum(a[i * K + k] * b[k * N + j] for k in range(K)) % mod for i in range(M) for j in range(N)]
        assert ints(C) == exp


def test_shape_ops_reference_vectors():
    """Oracle restatement of the shape / indexing kernels against the reference's unit-test vectors:
    host/mod.rs:613-663 (slice), :693-714 (diag), :717-761 (index_axis), :764-787 (transpose, concatenate),
    :1232-1288 (bit transposes), host/fixedpoint.rs:120-145 (fixed-point mean)."""
    x = mo.Ring.from_ints(64, [[1, 2], [3, 4]])
    assert mo.slice_ring([(1, None, None), (0, None, None)], x).to_ints().tolist() == [[3, 4]]
    assert mo.slice_ring([(-1, None, 2), (-1, None, 2)], x).to_ints().tolist() == [[4]]
    with pytest.raises(ValueError):
        mo.slice_ring([(-1, None, 2)] * 3, x)
    assert mo.diag(x).to_ints().tolist() == [1, 4]
    assert mo.diag(mo.Bits.from_array(np.array([[1, 1], [1, 0]]))).to_array().tolist() == [1, 0]
    x3 = mo.Ring.from_ints(64, [[[1, 2], [3, 4]], [[4, 5], [6, 7]]])
    assert mo.index_axis(0, 1, x3).to_ints().tolist() == [[4, 5], [6, 7]]
    v = mo.expand_dims([0], mo.index_axis(0, 1, mo.Ring.from_ints(64, [0, 1, 2, 3])))
    assert v.shape == (1,) and v.to_ints().tolist() == [1]
    b3 = mo.Bits.from_array(np.array([[[0, 1], [0, 0]], [[1, 1], [0, 0]]]))
    assert mo.index_axis(0, 1, b3).to_array().tolist() == [[1, 1], [0, 0]]
    assert mo.index_axis(0, 0, b3).to_array().tolist() == [[0, 1], [0, 0]]
    assert mo.transpose(x).to_ints().tolist() == [[1, 3], [2, 4]]
    c = mo.concat(0, [mo.expand_dims([0], x), mo.expand_dims([0], x)])
    assert c.to_ints().tolist() == [[[1, 2], [3, 4]], [[1, 2], [3, 4]]]
    xt = mo.transpose(mo.bit_decompose(mo.Ring.from_ints(64, [6743216615002642708])))
    assert xt.shape == (1, 64) and xt.to_array().reshape(-1).tolist() == [
        0, 0, 1, 0, 1, 0, 0, 0, 1, 0, 0, 1, 1, 0, 0, 1, 0, 1, 1, 1, 1, 1, 0, 1, 0, 0, 1, 0,
        0, 0, 1, 1, 0, 0, 0, 1, 0, 0, 1, 1, 1, 1, 1, 0, 1, 1, 0, 1, 0, 0, 1, 0, 1, 0, 0, 1,
        1, 0, 1, 1, 1, 0, 1, 0]
    assert mo.transpose(mo.Bits.from_array(np.array([[1, 1, 1], [0, 0, 0]]))).to_array().tolist() == [[1, 0]] * 3
    e = mo.fixed_encode(64, np.array([[1., 2.], [3., 4.]]), 16)
    assert mo.fixed_decode(mo.fixedpoint_mean(0, 2, 16, e), 32).tolist() == [2.0, 3.0]
    assert float(mo.fixed_decode(mo.fixedpoint_mean(None, 2, 16, e), 32)) == 2.5


def test_softmax_argmax_reference_vectors():
    """Oracle restatement of softmax / argmax against the reference's own test vectors and tolerances:
    replicated/softmax.rs:189-230 (abs error < 0.01), replicated/argmax.rs:123-200 (exact)."""
    s = mo.Session("kat", bytes(range(16)))
    x = np.array([1, 2.5, -3.0, 4.0, 2.0, -2.0, -2.0, -3.0, 3.0])
    e = np.exp(x - x.max())
    e /= e.sum()
    for bits, i_p, f_p in ((64, 8, 10), (128, 8, 27)):
        r = mo.rep_softmax(s, 0, 9, i_p, f_p, mo.rep_share(s, mo.fixed_encode(bits, x, f_p), 0))
        assert r.shape == (9,) and np.abs(mo.fixed_decode(mo.rep_reveal(r), f_p) - e).max() < 0.01
    cases = [(64, 0, 9, [1, 2, -3, 4, 2, 2, 2, 3, 105], 8),
             (64, 1, 4, [[1231, -323, -3, 12321], [93213, 12312321, -321, -3249], [3921, 4012, 3221, -321]], [3, 1, 1]),
             (128, 0, 9, [1, 2, -3, 10000, 10000, 10000, 10000, 10000, 10000], 3),
             (128, 1, 3, [[-3233, 329423, 54843], [3994, 123, -31326], [-324, 55321, 23441]], [1, 0, 1])]
    for bits, axis, upmost, vals, expect in cases:
        r = mo.rep_reveal(mo.rep_argmax(s, axis, upmost, mo.rep_share(s, mo.Ring.from_ints(bits, vals), 0)))
        assert r.bits == 64 and r.to_ints().tolist() == expect


def test_int2fl_equal_log_sqrt_reference_vectors():
    """replicated/log.rs:300-328 (test_int2fl: (1-2s)(1-z) v 2^p == x exactly), compare.rs:183-203 (test_equal),
    log.rs:330-432 (log2 / ln within 0.01 on Ring64 fixed(8,20), 0.01 / 0.001 on Ring128 fixed(10,30))."""
    s = mo.Session("kat", bytes(range(16)))
    x = mo.Ring.from_ints(64, [2, 10, 32, 0, -4])
    v, p, sg, z = (mo.rep_reveal(t).to_ints() for t in mo.int2fl(s, mo.rep_share(s, x, 0), 32, 5))
    sgn = lambda a: [int(t) - (1 << 64) if int(t) >> 63 else int(t) for t in a]
    v, p, sg, z = sgn(v), sgn(p), sgn(sg), sgn(z)
    assert [(1 - 2 * sg[i]) * (1 - z[i]) * v[i] * 2.0 ** p[i] for i in range(5)] == [2.0, 10.0, 32.0, 0.0, -4.0]
    a = mo.rep_share(s, mo.Ring.from_ints(64, [1024, 5, 4]), 0)
    b = mo.rep_share(s, mo.Ring.from_ints(64, [1024, 4, 5]), 1)
    assert mo.rep_reveal(mo.rep_equal(s, a, b)).to_array().tolist() == [1, 0, 0]
    xs = np.array([1.0, 2.0, 4.0, 8.0, 4.5, 10.5])
    X = mo.rep_share(s, mo.fixed_encode(64, xs, 20), 0)
    assert np.abs(mo.fixed_decode(mo.rep_reveal(mo.rep_log2(s, 8, 20, X)), 20) - np.log2(xs)).max() < 0.01
    xs = np.array([1.0, 2.5, 3.0, 4.0, 5.0])
    X = mo.rep_share(s, mo.fixed_encode(128, xs, 30), 0)
    assert np.abs(mo.fixed_decode(mo.rep_reveal(mo.rep_log(s, 10, 30, X)), 30) - np.log(xs)).max() < 0.001


def test_bit_and_ring_fill_sum_reference_vectors():
    """host/mod.rs:954-1009 (bit_fill, bit_xor, bit_or, bit_and), :1089-1116 (ring_fill, ring_sum with / without axis)."""
    assert mo.fill_bits(1, (2,)).to_array().tolist() == [1, 1]
    zero, one = mo.fill_bits(0, (5,)), mo.fill_bits(1, (5,))
    truth = {"xor": [one, one, zero, zero], "or": [one, one, one, zero], "and": [zero, zero, one, zero]}
    for op, want in truth.items():
        got = [mo.bit_op(op, a, b) for a, b in ((zero, one), (one, zero), (one, one), (zero, zero))]
        assert got == want, op
    assert ints(mo.fill(64, 1, (2,))) == [1, 1]
    x = mo.Ring.from_ints(64, [[1, 2], [3, 4]])
    assert ints(mo.sum_axis(x, 0)) == [4, 6]
    s = mo.sum_axis(x, None)
    assert s.shape == () and ints(s) == [10]
