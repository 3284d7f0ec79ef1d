"""The reference's own test vectors for the replicated protocols (moose/src/replicated/mod.rs tests), on the oracle:
TruncPr on powers of two and on negatives (mod.rs:909-960, result within one unit of the target, as the reference
asserts), msb (mod.rs:1002-1012), abs (mod.rs:1064-1074), ring_inject with and without a shift (mod.rs:1034-1058),
less / greater (mod.rs:1296-1360)."""
import numpy as np
import pytest

from oracle import moose_oracle as mo

KEY = bytes(range(16))
POWERS = [1 << i for i in range(63)]


def rep(sess, bits, values, owner=0):
    return mo.rep_share(sess, mo.Ring.from_ints(bits, values), owner)


def signed(r):
    m = 1 << r.bits
    return np.array([v - m if v >= m // 2 else v for v in r.to_ints().reshape(-1)], dtype=object)


@pytest.mark.parametrize("bits", [64, 128])
@pytest.mark.parametrize("values,amount,target", [
    (POWERS, 1, [0] + POWERS[:62]),
    (POWERS, 62, [0] * 62 + [1]),
    (POWERS, 61, [0] * 61 + [1, 2]),
    ([-10], 1, [-5]),
    ([-10], 0, [-10]),
    ([-1152921504606846976], 60, [-1]),
])
def test_trunc_pr_reference_cases(bits, values, amount, target):
    sess = mo.Session("vectors", KEY)
    x = rep(sess, bits, values, owner=2)                     # shared by carole, opened at carole (mod.rs:882-886)
    got = signed(mo.rep_reveal(mo.rep_trunc_pr(sess, amount, x), 2))
    diff = np.array(target, dtype=object) - got
    assert all(d in (-1, 0, 1) for d in diff), diff


@pytest.mark.parametrize("bits", [64, 128])
def test_msb_abs_reference_cases(bits):
    sess = mo.Session("vectors", KEY)
    x = rep(sess, bits, [-10, -100, -200000, 0, 1])
    assert list(mo.rep_reveal(mo.rep_msb_ring(sess, x)).to_ints()) == [1, 1, 1, 0, 0]
    y = rep(sess, bits, [-10, -100, -200000, 0, 1000])
    assert list(mo.rep_reveal(mo.rep_abs(sess, y)).to_ints()) == [10, 100, 200000, 0, 1000]


@pytest.mark.parametrize("xs", [[0, 1, 1, 0, 0], [[0, 1], [1, 0]]])
def test_ring_inject_reference_cases(xs):
    sess = mo.Session("vectors", KEY)
    b = mo.rep_share(sess, mo.Bits.from_array(np.array(xs, dtype=np.uint8)), 0)
    want = np.array(xs, dtype=object)
    for bits in (64, 128):
        got = mo.rep_reveal(mo.rep_ring_inject(sess, bits, 0, b)).to_ints()
        assert np.array_equal(got, want)
    got = mo.rep_reveal(mo.rep_ring_inject(sess, 64, 20, b)).to_ints()
    assert np.array_equal(got, want * (1 << 20))


@pytest.mark.parametrize("bits", [64, 128])
def test_less_greater_reference_cases(bits):
    sess = mo.Session("vectors", KEY)
    x, y = [0, 1, 2, -1, -2], [-1, -2, 3, -1, -1]
    lt = mo.rep_reveal(mo.rep_less(sess, rep(sess, bits, x), rep(sess, bits, y)))
    assert list(lt.to_array()) == [0, 0, 1, 0, 1]
    if bits == 64:
        x, y = x + [1 << 62], y + [-(1 << 62) + 1]
    gt = mo.rep_reveal(mo.rep_greater(sess, rep(sess, bits, x), rep(sess, bits, y)))
    assert list(gt.to_array()) == [1, 1, 0, 0, 0] + ([1] if bits == 64 else [])


# ---- fixed-point math: the vectors and tolerances of moose/src/fixedpoint/ops.rs tests ----------------------------------
def fixed(sess, bits, frac, values, owner=0):
    return mo.rep_share(sess, mo.fixed_encode(bits, np.asarray(values, dtype=np.float64), frac), owner)


def decoded(x, frac):
    return mo.fixed_decode(mo.rep_reveal(x), frac)


@pytest.mark.parametrize("bits,i_p,f_p", [(64, 10, 15), (128, 10, 15)])
def test_division_reference_vectors(bits, i_p, f_p):
    """fixedpoint/ops.rs:3126-3145: [1, 2, 2] / [3, 7, 1.41], squared distance below 2^-15."""
    sess = mo.Session("vectors", KEY)
    x, y = np.array([1.0, 2.0, 2.0]), np.array([3.0, 7.0, 1.41])
    z = mo.rep_div(sess, i_p, f_p, fixed(sess, bits, f_p, x), fixed(sess, bits, f_p, y))
    assert ((decoded(z, f_p) - x / y) ** 2).max() < 2.0 ** -15


@pytest.mark.parametrize("fn,ref,bits,i_p,f_p,tol,xs", [
    (mo.rep_pow2, np.exp2, 64, 10, 10, 1e-3, [1, 2.5, -3.0, 4.0]),                      # ops.rs:3408-3420
    (mo.rep_pow2, np.exp2, 128, 30, 10, 1e-3, [1, 2.5, -3.0, 4.0]),
    (mo.rep_exp, np.exp, 64, 10, 10, 1e-1, [1, 2.5, -3.0, 4.0]),                        # ops.rs:3422-3434
    (mo.rep_exp, np.exp, 128, 20, 20, 1e-3, [1, 2.5, -3.0, 4.0]),
    (mo.rep_sqrt, np.sqrt, 64, 10, 10, 1e-1, [0.001, 0.01, 0.1, 1, 2, 3, 4, 10, 20, 30, 40, 50, 100, 1000]),   # :3436-3455
    (mo.rep_sqrt, np.sqrt, 128, 20, 20, 1e-3, [0.001, 0.01, 0.1, 1, 2, 3, 4, 10, 50, 100, 1000, 10000, 100000, 500000]),
    (mo.rep_sigmoid, lambda v: 1 / (1 + np.exp(-v)), 64, 10, 10, 1e-1, [1, 2.5, -3.0, 4.0, 40.0, -30.0, -4.0, -6.0, 6.0]),
    (mo.rep_sigmoid, lambda v: 1 / (1 + np.exp(-v)), 128, 20, 20, 1e-3, [1, 2.5, -3.0, 4.0, -4.0, -6.0, 6.0, 1024.0]),
])
def test_unary_math_reference_vectors(fn, ref, bits, i_p, f_p, tol, xs):
    """rep_approx_unary_fixed_test (ops.rs:3362-3406): x is a host fixed tensor at alice, the result is opened at alice."""
    sess = mo.Session("vectors", KEY)
    x = np.array(xs, dtype=np.float64)
    got = decoded(fn(sess, i_p, f_p, fixed(sess, bits, f_p, x)), f_p)
    assert np.abs(got - ref(x)).max() < tol


@pytest.mark.parametrize("bits", [64, 128])
def test_prefix_mul_and_polynomial_reference_vectors(bits):
    """prefix_mul of [1, 2, 3, 4] at 15 fractional bits -> 1, 2, 6, 24 (ops.rs:3283-3310); 1 + 2x + 3x^2 at
    x = 1..4 -> 6, 17, 34, 57 (ops.rs:3341-3358)."""
    sess = mo.Session("vectors", KEY)
    f = 15
    xs = [fixed(sess, bits, f, [v]) for v in (1.0, 2.0, 3.0, 4.0)]
    pre = mo.prefix_op(sess, xs, lambda s, a, b: mo.rep_trunc_pr(s, f, mo.rep_mul(s, a, b)))
    assert [int(decoded(p, f)[0]) for p in pre] == [1, 2, 6, 24]
    y = mo.polynomial_eval(sess, f, [1.0, 2.0, 3.0], fixed(sess, bits, f, [1.0, 2.0, 3.0, 4.0]))
    assert np.abs(decoded(y, f) - np.array([6.0, 17.0, 34.0, 57.0])).max() < 2.0 ** -13   # a few truncation units


@pytest.mark.parametrize("bits", [64, 128])
def test_trunc_pr_and_msb_fuzz_like_the_reference(bits):
    """test_fuzzy_rep_trunc64/128 (mod.rs:960-978; inputs (x >> 2) - 1, amount < bits - 2, result within one unit of
    x >> amount), test_fuzzy_adt_trunc64/128 (additive/trunc.rs:272-286: the same bound on the two-party form) and
    test_fuzzy_rep_msb64/128 (mod.rs:1014-1030)."""
    rng = np.random.default_rng(bits)
    sess = mo.Session("vectors", KEY)
    mod = 1 << bits
    for _ in range(12):
        n = int(rng.integers(1, 5))
        raw = [((int.from_bytes(rng.bytes(bits // 8), "little") >> 2) - 1) % mod for _ in range(n)]
        amount = int(rng.integers(0, bits - 2))
        x = rep(sess, bits, raw, owner=2)
        got = mo.rep_reveal(mo.rep_trunc_pr(sess, amount, x), 2).to_ints().reshape(-1)
        assert all((int(g) - (v >> amount)) % mod in (0, 1, mod - 1) for g, v in zip(got, raw)), (raw, amount)
        adt = mo.adt_trunc_pr(sess, amount, mo.rep_to_adt(x))
        got = mo.adt_reveal(adt).to_ints().reshape(-1)
        assert all((int(g) - (v >> amount)) % mod in (0, 1, mod - 1) for g, v in zip(got, raw)), (raw, amount)
        full = [int.from_bytes(rng.bytes(bits // 8), "little") for _ in range(n)]
        msb = mo.rep_reveal(mo.rep_msb_ring(sess, rep(sess, bits, full))).to_ints().reshape(-1)
        assert [int(m) for m in msb] == [v >> (bits - 1) for v in full]
