"""Replicated fixed-point math on the LogReg path -- division, pow2, exp, sigmoid -- through the C ABI:
share-level parity with the oracle's op-for-op restatement, and the absolute-error tolerances of the
reference's own tests (moose/src/fixedpoint/ops.rs:3360-3468: 1e-3 on Ring128 with fixed(20,20),
1e-1 on Ring64 with fixed(10,10))."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import RING64, RING128
from oracle import moose_oracle as mo
from _util import ENTROPY_KEY, SESSION_ID, same_rep

pytestmark = pytest.mark.gpu
DT = {64: RING64, 128: RING128}


def pair():
    return mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)


def shared(g, o, bits, frac, vals, owner=0):
    v = np.asarray(vals, dtype=np.float64)
    gx = g.share(g.fixedpoint_encode(DT[bits], frac, g.from_host(mb.FLOAT64, v)), owner)
    ox = mo.rep_share(o, mo.fixed_encode(bits, v, frac), owner)
    return gx, ox


def decode(g, frac, rep):
    return g.fixedpoint_decode(frac, g.reveal(rep)).to_host()


@pytest.mark.parametrize("bits", [64, 128])
def test_division_reference_vectors(bits):
    """fixedpoint/ops.rs:3126-3145 (test_fixed_rep_div64/128): [1,2,2] / [3,7,1.41] at fixed(10,15),
    squared distance below 2^-15."""
    g, o = pair()
    x, y = np.array([1.0, 2.0, 2.0]), np.array([3.0, 7.0, 1.41])
    gx, ox = shared(g, o, bits, 15, x)
    gy, oy = shared(g, o, bits, 15, y)
    gz, oz = g.rep_div(10, 15, gx, gy), mo.rep_div(o, 10, 15, ox, oy)
    assert same_rep(gz, oz)
    assert ((decode(g, 15, gz) - x / y) ** 2).max() < 2.0 ** -15
    g.close()


@pytest.mark.parametrize("bits,i_p,f_p,tol", [(128, 20, 20, 1e-3), (64, 10, 10, 1e-1), (128, 24, 40, 5e-2)])
def test_division(bits, i_p, f_p, tol):
    """Tolerances: the reference's own (1e-3 / 1e-1).  At fixed(24,40) -- pymoose's default dtype -- the Goldschmidt
    step a*a carries 4f = 160 fractional bits and wraps in Ring128 (division.rs:67-73), which costs the protocol
    itself ~1e-2 of accuracy; shares still match the oracle bit for bit.  Positive divisors only: the reference's norm() up-shifts x instead of |x| (division.rs:128), so its
    Goldschmidt start value is wrong for negative divisors; this port reproduces the protocol as written."""
    g, o = pair()
    x = np.array([[1.0, 2.0, -3.5, 0.25], [10.0, -0.75, 7.0, 1.0]])
    y = np.array([[2.0, 4.0, 1.25, 0.5], [3.0, 0.6, 7.0, 9.0]])
    gx, ox = shared(g, o, bits, f_p, x)
    gy, oy = shared(g, o, bits, f_p, y, owner=1)
    gz, oz = g.rep_div(i_p, f_p, gx, gy), mo.rep_div(o, i_p, f_p, ox, oy)
    assert same_rep(gz, oz)
    assert np.abs(decode(g, f_p, gz) - x / y).max() < tol
    g.close()


def test_division_precision_guard():
    """division.rs:33-38: 2*(i+f) must fit the ring (fixed(14,23) on Ring64 is rejected)."""
    g, _ = pair()
    x = g.share(g.fill(RING64, 1 << 23, (4,)), 0)
    with pytest.raises(mb.MooseB200Error) as e:
        g.rep_div(14, 23, x, x)
    assert "BitLength = 64" in str(e.value)
    g.close()


@pytest.mark.parametrize("bits,i_p,f_p,tol", [(128, 20, 20, 1e-3), (128, 24, 40, 1e-2)])
def test_exp_pow2_sigmoid(bits, i_p, f_p, tol):
    g, o = pair()
    xs = np.array([-6.0, -2.5, -1.0, -0.1, 0.0, 0.3, 1.0, 2.0, 4.0, 6.5])
    gx, ox = shared(g, o, bits, f_p, xs)
    ge, oe = g.rep_exp(i_p, f_p, gx), mo.rep_exp(o, i_p, f_p, ox)
    assert same_rep(ge, oe)
    assert (np.abs(decode(g, f_p, ge) - np.exp(xs)) / np.maximum(np.exp(xs), 1.0)).max() < tol
    gp, op = g.rep_pow2(i_p, f_p, gx), mo.rep_pow2(o, i_p, f_p, ox)
    assert same_rep(gp, op)
    assert (np.abs(decode(g, f_p, gp) - 2.0 ** xs) / np.maximum(2.0 ** xs, 1.0)).max() < tol
    gs, os_ = g.sigmoid(i_p, f_p, gx), mo.rep_sigmoid(o, i_p, f_p, ox)
    assert same_rep(gs, os_)
    assert np.abs(decode(g, f_p, gs) - 1.0 / (1.0 + np.exp(-xs))).max() < tol
    g.close()


def test_sigmoid_saturation():
    """arith.rs:717-741: inputs beyond +-ln 2^(i-1) are clamped to exactly 1 / 0."""
    g, o = pair()
    i_p, f_p = 8, 27
    xs = np.array([-20.0, -6.0, 0.0, 6.0, 20.0, 100.0, -100.0, 3.0])
    gx, ox = shared(g, o, 128, f_p, xs)
    gs, os_ = g.sigmoid(i_p, f_p, gx), mo.rep_sigmoid(o, i_p, f_p, ox)
    assert same_rep(gs, os_)
    out = decode(g, f_p, gs)
    lim = np.log(2.0 ** (i_p - 1))
    assert np.all(out[xs > lim] == 1.0) and np.all(out[xs < -lim] == 0.0)
    inside = np.abs(xs) <= lim
    assert np.abs(out[inside] - 1.0 / (1.0 + np.exp(-xs[inside]))).max() < 1e-3
    g.close()


def test_greater_and_fill():
    g, o = pair()
    a = np.array([1.0, -2.0, 3.0, 0.5, -0.5, 7.0])
    b = np.array([0.5, -1.0, 3.0, 2.0, -0.75, -7.0])
    gx, ox = shared(g, o, 64, 23, a)
    gy, oy = shared(g, o, 64, 23, b, owner=2)
    gg, og = g.rep_greater(gx, gy), mo.rep_greater(o, ox, oy)
    assert same_rep(gg, og)
    assert g.reveal(gg).bits().tolist() == [int(p > q) for p, q in zip(a, b)]
    gf, of = g.rep_fill(RING128, 12345, (2, 3)), mo.rep_fill(128, 12345, (2, 3))
    assert same_rep(gf, of)
    g.close()


@pytest.mark.parametrize("bits,i_p,f_p,shape", [(128, 24, 40, (64, 2)), (128, 20, 20, (256,)), (64, 10, 15, (128,)), (128, 8, 27, (3, 128))])
def test_batched_math_paths_share_parity(bits, i_p, f_p, shape):
    """Element counts that are multiples of 128 take the batched kernels of csrc/math_fused.cu (prefix-or levels,
    inject + sum, Pow2FromBits in registers) instead of one launch per replicated op: same shares as the oracle's
    op-by-op protocol, bit for bit, for division (top_most_index), pow2 / exp (higher bits) and sigmoid."""
    g, o = pair()
    rng = np.random.default_rng(hash((bits, i_p, f_p, shape)) % 2 ** 32)
    lim = min(6.0, 0.6 * (i_p - 1))
    xs = rng.uniform(-lim, lim, shape)
    ys = rng.uniform(0.3, 2.0 ** min(i_p - 2, 5), shape)
    gx, ox = shared(g, o, bits, f_p, xs)
    gy, oy = shared(g, o, bits, f_p, ys, owner=1)
    l0 = g.launch_count
    gz, oz = g.rep_div(i_p, f_p, gx, gy), mo.rep_div(o, i_p, f_p, ox, oy)
    div_launches = g.launch_count - l0
    assert same_rep(gz, oz)
    assert div_launches < 120, div_launches          # ~1000 when every replicated op is its own launch
    if bits == 128:   # pow2 / exp / sigmoid need 2 (i + f) <= ring bits for their inner division
        gp, op = g.rep_pow2(i_p, f_p, gx), mo.rep_pow2(o, i_p, f_p, ox)
        assert same_rep(gp, op)
        l0 = g.launch_count
        gs, os_ = g.sigmoid(i_p, f_p, gx), mo.rep_sigmoid(o, i_p, f_p, ox)
        sig_launches = g.launch_count - l0
        assert same_rep(gs, os_)
        assert sig_launches < 600, sig_launches
        assert np.abs(decode(g, f_p, gs) - 1.0 / (1.0 + np.exp(-xs))).max() < 2e-2
    g.close()


def test_batched_and_op_by_op_paths_agree():
    """A/B inside the library: the same call with the batched paths switched off (fused bit kernels off selects the
    op-by-op composition everywhere) yields the same shares."""
    xs = np.random.default_rng(11).uniform(-4, 4, (128,))
    outs = []
    for fused in (True, False):
        g, _ = pair()
        g.set_fused_bits(fused)
        gx = g.share(g.fixedpoint_encode(RING128, 27, g.from_host(mb.FLOAT64, xs)), 0)
        s = g.sigmoid(8, 27, gx)
        outs.append([s.share(i).to_host() for i in range(3)])
        g.close()
    for i in range(3):
        assert np.array_equal(outs[0][i], outs[1][i])


@pytest.mark.parametrize("bits,i_p,f_p", [(64, 8, 10), (128, 24, 40)])
def test_math_on_zero_element_tensors(bits, i_p, f_p):
    """Division, exp and sigmoid of an empty batch: empty results, the reference's entropy requests still drawn (a
    non-empty sigmoid after them equals the oracle's share by share)."""
    g, o = pair()
    empty = np.zeros((0, 2))
    gx, ox = shared(g, o, bits, f_p, empty, 0)
    gy, oy = shared(g, o, bits, f_p, empty, 1)
    for gr, orr in ((g.rep_div(i_p, f_p, gx, gy), mo.rep_div(o, i_p, f_p, ox, oy)),
                    (g.rep_exp(i_p, f_p, gx), mo.rep_exp(o, i_p, f_p, ox)),
                    (g.sigmoid(i_p, f_p, gx), mo.rep_sigmoid(o, i_p, f_p, ox))):
        assert same_rep(gr, orr) and g.reveal(gr).shape == (0, 2)
    gz, oz = shared(g, o, bits, f_p, [[0.5, -1.25]], 2)
    assert same_rep(g.sigmoid(i_p, f_p, gz), mo.rep_sigmoid(o, i_p, f_p, oz))
    g.close()
