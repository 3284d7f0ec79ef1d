"""More of the reference's own known answers for the replicated layer, on the oracle (CPU):

  * rep add / mul / dot on Ring64 and Ring128, literal cases and the wrap-around property the reference fuzzes
    (moose/src/replicated/mod.rs:554-696);
  * mirrored (public) add / sub / mul, both operand orders (mod.rs:698-826);
  * bit_decompose of 2^30 - 1 shared by alice and by bob, bit_compose(bit_decompose(x)) = x (mod.rs:1089-1224);
  * prefix_or / prefix_and over the 64 planes of one element (mod.rs:1226-1274);
  * binary_adder on the reference's two 64-bit operands, norm(896) = (3584, 4), approximate_reciprocal(3.5) ~ 74
    (replicated/division.rs:257-324);
  * pow2_from_bits([0,1,1,1]) = 2^14 (replicated/exp.rs:253-274);
  * mux with a ring selector and with a bit selector (replicated/control_flow.rs:53-123);
  * additive sharing arithmetic as the lowered graph performs it: share-wise add, public operand on the share of
    the host that holds it (additive/ops.rs:296-337).
"""
import numpy as np
import pytest

from oracle import moose_oracle as mo

KEY = bytes(range(16))
M64, M128 = (1 << 64) - 1, (1 << 128) - 1


def ring(bits, values, shape=None):
    return mo.Ring.from_ints(bits, values, shape)


def ints(r):
    return [int(v) for v in r.to_ints().reshape(-1)]


def shared(sess, bits, values, owner=0, shape=None):
    return mo.rep_share(sess, ring(bits, values, shape), owner)


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("x,y,z", [([1, 2, 3], [1, 2, 3], [2, 4, 6]), ([-1, -2, -3], [1, 2, 3], [0, 0, 0])])
def test_rep_add_reference_cases(bits, x, y, z):
    sess = mo.Session("vectors", KEY)
    got = mo.rep_reveal(mo.rep_add(shared(sess, bits, x), shared(sess, bits, y)))
    assert ints(got) == z


@pytest.mark.parametrize("bits", [64, 128])
def test_rep_mul_and_dot_wrap_like_the_ring(bits):
    """test_fuzzy_rep_mul64/128 and test_fuzzy_rep_dot64/128: full-range operands of length 1..24, every receiver."""
    rng = np.random.default_rng(bits)
    mask = (1 << bits) - 1
    sess = mo.Session("vectors", KEY)
    for length in (1, 2, 7, 24):
        a = [int.from_bytes(rng.bytes(bits // 8), "little") for _ in range(length)]
        b = [int.from_bytes(rng.bytes(bits // 8), "little") for _ in range(length)]
        xs, ys = shared(sess, bits, a, owner=length % 3), shared(sess, bits, b, owner=(length + 1) % 3)
        prod = mo.rep_mul(sess, xs, ys)
        dot = mo.rep_dot(sess, xs, ys)
        for receiver in range(3):
            assert ints(mo.rep_reveal(prod, receiver)) == [(p * q) & mask for p, q in zip(a, b)]
            assert ints(mo.rep_reveal(dot, receiver)) == [sum(p * q for p, q in zip(a, b)) & mask]


@pytest.mark.parametrize("bits,mask", [(64, M64), (128, M128)])
def test_mirrored_add_sub_mul_reference_cases(bits, mask):
    sess = mo.Session("vectors", KEY)
    two = mo.mir_fill(bits, 2, (3,))
    x = shared(sess, bits, [0, 1, 2])
    assert ints(mo.rep_reveal(mo.rep_add_mir(x, two))) == [2, 3, 4]
    assert ints(mo.rep_reveal(mo.rep_mir_add(two, x))) == [2, 3, 4]
    y = shared(sess, bits, [2, 3, 4])
    assert ints(mo.rep_reveal(mo.rep_sub_mir(y, two))) == [0, 1, 2]
    assert ints(mo.rep_reveal(mo.rep_mir_sub(two, y))) == [0, mask, mask - 1]
    # the reference multiplies by a mirrored value of shape [1] (mod.rs:784-796)
    two1 = mo.mir_fill(bits, 2, (1,))
    assert ints(mo.rep_reveal(mo.rep_mul_mir(x, two1))) == [0, 2, 4]
    assert ints(mo.rep_reveal(mo.rep_mir_mul(two1, x))) == [0, 2, 4]


@pytest.mark.parametrize("owner", [0, 1])
def test_bit_decompose_reference_case(owner):
    sess = mo.Session("vectors", KEY)
    planes = mo.rep_reveal(mo.rep_bit_decompose(sess, shared(sess, 64, [1073741823], owner=owner)))
    assert planes.shape == (64, 1)
    assert list(planes.to_array().reshape(-1)) == [1] * 30 + [0] * 34


@pytest.mark.parametrize("bits", [64, 128])
def test_bit_compose_inverts_bit_decompose(bits):
    sess = mo.Session("vectors", KEY)
    x = shared(sess, bits, [1073741823, 0, 6])
    back = mo.rep_bit_compose(sess, bits, mo.rep_bit_decompose(sess, x))
    assert ints(mo.rep_reveal(back)) == [1073741823, 0, 6]


@pytest.mark.parametrize("fn,value,target", [
    (mo.prefix_or, 1024, [0] * 10 + [1] * 54),
    (mo.prefix_and, 7, [1] * 3 + [0] * 61),
])
def test_prefix_or_and_reference_cases(fn, value, target):
    sess = mo.Session("vectors", KEY)
    planes = mo.rep_bit_decompose(sess, shared(sess, 64, [value]))
    out = fn(sess, [mo.rep_index(i, planes) for i in range(64)])
    assert [int(mo.rep_reveal(el).to_array().reshape(-1)[0]) for el in out] == target


def test_binary_adder_reference_case():
    sess = mo.Session("vectors", KEY)
    x, y = ring(64, [3884509700957842751]), ring(64, [13611438098135434720])
    want = mo.bit_decompose(mo.add(x, y))
    xs = mo.rep_share(sess, mo.bit_decompose(x), 0)
    ys = mo.rep_share(sess, mo.bit_decompose(y), 0)
    assert mo.rep_reveal(mo.binary_adder(sess, xs, ys, 64)) == want


def test_norm_and_reciprocal_reference_cases():
    sess = mo.Session("vectors", KEY)
    x = shared(sess, 64, [896])                       # 3.5 * 2^8
    upshifted, topmost = mo.div_norm(sess, 64, 12, x)
    assert ints(mo.rep_reveal(topmost)) == [4]
    assert ints(mo.rep_reveal(upshifted)) == [3584]
    approx = ints(mo.rep_reveal(mo.approximate_reciprocal(sess, 64, 4, 8, x)))[0]
    assert abs(approx - 74) <= 1


def test_pow2_from_bits_reference_case():
    sess = mo.Session("vectors", KEY)
    x = shared(sess, 64, [0, 1, 1, 1], shape=(4, 1))
    bits = [mo.rep_index_axis(0, i, x) for i in range(4)]
    assert ints(mo.rep_reveal(mo.pow2_from_bits(sess, 64, bits))) == [16384]


def test_mux_reference_cases():
    sess = mo.Session("vectors", KEY)
    enc = lambda v: mo.fixed_encode(128, np.array(v, dtype=np.float64), 24)
    a = mo.rep_share(sess, enc([1.0, 1.0, -1.0]), 0)
    x = mo.rep_share(sess, enc([1.0, 2.0, 3.0]), 0)
    y = mo.rep_share(sess, enc([4.0, 5.0, 6.0]), 1)
    s = mo.rep_sub(mo.rep_fill(128, 1, (3,)), mo.rep_msb_ring(sess, a))
    got = mo.fixed_decode(mo.rep_reveal(mo.rep_mux(sess, s, x, y)), 24)
    assert list(got) == [1.0, 2.0, 6.0]

    sel = mo.rep_share(sess, mo.Bits.from_array(np.array([1, 0, 1, 0], dtype=np.uint8)), 0)
    x = mo.rep_share(sess, enc([1.0, 2.0, 3.0, 4.0]), 0)
    y = mo.rep_share(sess, enc([4.0, 5.0, 6.0, -1.0]), 1)
    got = mo.fixed_decode(mo.rep_reveal(mo.rep_mux_bit_selector(sess, 128, sel, x, y)), 24)
    assert list(got) == [1.0, 5.0, 3.0, -1.0]


def test_additive_sharing_arithmetic_reference_case():
    """adt + adt is share-wise; adt + host adds the public tensor to the share held by that host."""
    x = mo.AdtTensor([ring(64, [1, 2, 3]), ring(64, [4, 5, 6])])
    y = mo.AdtTensor([ring(64, [7, 8, 9]), ring(64, [1, 2, 3])])
    z = [mo.add(a, b) for a, b in zip(x.shares, y.shares)]
    assert ints(z[0]) == [8, 10, 12] and ints(z[1]) == [5, 7, 9]
    r = ring(64, [7, 8, 9])
    assert ints(mo.add(x.shares[0], r)) == [8, 10, 12] and ints(x.shares[1]) == [4, 5, 6]   # r at alice
    assert ints(x.shares[0]) == [1, 2, 3] and ints(mo.add(x.shares[1], r)) == [11, 13, 15]  # r at bob
    assert ints(mo.adt_reveal(mo.AdtTensor(z))) == [13, 17, 21]


def test_identity_mean_add_n_sum_diag_index_reference_cases():
    """replicated/mod.rs:281-292 (share / reveal identity), :371-391 (mean_as_fixedpoint + trunc_pr -> 2.0 within 0.01),
    :394-423 (add_n over three owners, Ring64 and Ring128), :426-437 (sum without axis), :441-478 (diag on ring and bit
    tensors), :480-515 (index_axis 0, 1), :517-552 (index of a replicated bit array)."""
    sess = mo.Session("vectors", KEY)
    assert ints(mo.rep_reveal(shared(sess, 64, [1, 2, 3]))) == [1, 2, 3]
    x = mo.rep_share(sess, mo.fixed_encode(64, np.array([1.0, 2.0, 3.0]), 24), 0)
    mean = mo.rep_trunc_pr(sess, 24, mo.rep_fixedpoint_mean(None, 2, 24, x))
    assert abs(2.0 - float(mo.fixed_decode(mo.rep_reveal(mean), 24))) < 0.01
    parts = [shared(sess, 64, [1, 2, 3], 0), shared(sess, 64, [2, 3, 4], 1), shared(sess, 64, [5, 12, 13], 2)]
    assert ints(mo.rep_reveal(mo.rep_add_n(parts))) == [8, 17, 20]
    parts = [shared(sess, 128, [1, 2, 3, 2, 3, 4], 0, (2, 3)), shared(sess, 128, [2, 3, 4, 2, 3, 4], 1, (2, 3)),
             shared(sess, 128, [5, 12, 13, 1, 2, 3], 2, (2, 3))]
    total = mo.rep_reveal(mo.rep_add_n(parts))
    assert total.shape == (2, 3) and ints(total) == [8, 17, 20, 5, 8, 11]
    s = mo.rep_reveal(mo.rep_sum(None, shared(sess, 64, [1, 2, 3])))
    assert s.shape == () and ints(s) == [6]
    for bits in (64, 128):
        m = shared(sess, bits, [1, 2, 3, 4], shape=(2, 2))
        d = mo.RepTensor([[mo.diag(v) for v in row] for row in m.shares])
        assert ints(mo.rep_reveal(d)) == [1, 4]
        cube = shared(sess, bits, [1, 2, 3, 4, 4, 5, 6, 7], shape=(2, 2, 2))
        got = mo.rep_reveal(mo.rep_index_axis(0, 1, cube))
        assert got.shape == (2, 2) and ints(got) == [4, 5, 6, 7]
    b = mo.rep_share(sess, mo.Bits.from_array(np.array([[1, 0], [1, 1]], dtype=np.uint8)), 0)
    d = mo.RepTensor([[mo.diag(v) for v in row] for row in b.shares])
    assert mo.rep_reveal(d).to_array().tolist() == [1, 1]
    assert mo.rep_reveal(mo.rep_index(0, b)).to_array().tolist() == [1, 0]
    assert mo.rep_reveal(mo.rep_index_axis(0, 1, b)).to_array().tolist() == [1, 1]
