"""Predictor-level parity (BASELINE configs 3/4 at oracle-sized batches): the MLP op sequence of the pymoose
MLPRegressor on Ring64 fixed(14,23), shares compared bit for bit with the oracle under the shared entropy
source, outputs compared with a float64 numpy MLP at the tolerance the reference's own predictor tests use
(decimal=1 for MLP regressors, multilayer_perceptron_predictor_test.py:100-206)."""
import numpy as np
import pytest

import moose_b200 as mb
from moose_b200.predictors import LinearPredictor, MLPPredictor
from oracle import moose_oracle as mo
from oracle import predictors_oracle as po
from _util import ENTROPY_KEY, SESSION_ID, same_rep

pytestmark = pytest.mark.gpu


def make_mlp(rng, dims):
    ws = [rng.normal(0, dims[i] ** -0.5, (dims[i], dims[i + 1])) for i in range(len(dims) - 1)]
    bs = [rng.normal(0, 0.1, dims[i + 1]) for i in range(len(dims) - 1)]
    return ws, bs


def numpy_mlp(x, ws, bs):
    for i, (w, b) in enumerate(zip(ws, bs)):
        x = x @ w + b
        if i < len(ws) - 1:
            x = np.maximum(x, 0)
    return x


@pytest.mark.parametrize("dims,batch", [((32, 16, 16, 4), 24), ((784, 512, 512, 10), 128)])
@pytest.mark.parametrize("relu_via", ["maximum", "relu"])
def test_mlp_ring64_share_parity(dims, batch, relu_via):
    if relu_via == "relu" and dims[0] == 784:
        pytest.skip("one large case is enough")
    rng = np.random.default_rng(3)
    ws, bs = make_mlp(rng, dims)
    x = rng.uniform(0, 1, (batch, dims[0]))
    g, o = mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)
    try:
        pred = MLPPredictor(g, ws, bs, mb.RING64, 23, relu_via)
        gout = pred.forward_shared(pred.encode(x))
        oout, odec = po.mlp(o, x, ws, bs, 64, 23, relu_via)
        assert same_rep(gout, oout)
        dec = g.fixedpoint_decode(23, g.reveal(gout)).to_host()
        assert np.array_equal(dec, odec)
        np.testing.assert_almost_equal(dec, numpy_mlp(x, ws, bs), decimal=1)
        assert np.abs(dec - numpy_mlp(x, ws, bs)).max() < 1e-3
    finally:
        g.close()


def test_mlp_bias_added_as_a_public_value():
    """bias_via="public": the mirrored bias is added to share 0 only (the concrete rep + mirrored kernel,
    replicated/arith.rs:81-121) instead of going through ShareOp at player 0 as in the logical graph (the default,
    fixedpoint/ops.rs:30-52, covered by test_mlp_ring64_share_parity); same values, share-level parity with the oracle."""
    rng = np.random.default_rng(8)
    dims, batch = (32, 16, 16, 4), 24
    ws, bs = make_mlp(rng, dims)
    x = rng.uniform(0, 1, (batch, dims[0]))
    g, o = mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)
    try:
        pred = MLPPredictor(g, ws, bs, mb.RING64, 23, "maximum", bias_via="public")
        gout = pred.forward_shared(pred.encode(x))
        oout, odec = po.mlp(o, x, ws, bs, 64, 23, "maximum", bias_via="public")
        assert same_rep(gout, oout)
        dec = g.fixedpoint_decode(23, g.reveal(gout)).to_host()
        assert np.array_equal(dec, odec)
        assert np.abs(dec - numpy_mlp(x, ws, bs)).max() < 1e-3
    finally:
        g.close()


def test_linear_ring128_dot_trunc_prefix():
    rng = np.random.default_rng(2)
    feats, batch = 512, 300
    coeffs, intercepts = rng.normal(0, feats ** -0.5, (2, feats)), rng.normal(0, 0.1, (1, 2))
    x = rng.uniform(-1, 1, (batch, feats))
    g = mb.Session(0, SESSION_ID, ENTROPY_KEY)
    try:
        out = LinearPredictor(g, coeffs, intercepts)(x)
        np.testing.assert_almost_equal(out, x @ coeffs.T + intercepts, decimal=5)
    finally:
        g.close()


def test_logreg_sigmoid_ring128_share_parity():
    """BASELINE config 3 at an oracle-sized batch: LinearClassifier(SIGMOID, 2 classes), 512 features, Ring128
    fixed(24,40); classifier outputs are checked at decimal=2 like the reference's predictor tests
    (linear_predictor_test.py:81-115)."""
    rng = np.random.default_rng(1)
    feats, batch = 512, 64
    coeffs, intercepts = rng.normal(0, feats ** -0.5, (2, feats)), rng.normal(0, 0.1, (1, 2))
    x = rng.uniform(-1, 1, (batch, feats))
    g, o = mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)
    try:
        pred = LinearPredictor(g, coeffs, intercepts, post_transform="sigmoid")
        gout = pred.forward_shared(pred.encode(x))
        oout, odec = po.linear(o, x, coeffs, intercepts, post_transform="sigmoid")
        assert same_rep(gout, oout)
        dec = g.fixedpoint_decode(40, g.reveal(gout)).to_host()
        ref = 1.0 / (1.0 + np.exp(-(x @ coeffs.T + intercepts)))
        np.testing.assert_almost_equal(dec, ref, decimal=2)
    finally:
        g.close()


@pytest.mark.parametrize("post,n_classes", [("softmax", 3), ("normalized_sigmoid", 3)])
def test_linear_classifier_multiclass_share_parity(post, n_classes):
    """LinearClassifier with SOFTMAX / multi-class SIGMOID (linear_predictor.py:160-175, 294-301), Ring128 fixed(24,40);
    decimal=2 like the reference's predictor tests."""
    rng = np.random.default_rng(11)
    feats, batch = 16, 24
    coeffs, intercepts = rng.normal(0, 0.5, (n_classes, feats)), rng.normal(0, 0.1, (1, n_classes))
    x = rng.uniform(-1, 1, (batch, feats))
    g, o = mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)
    try:
        pred = LinearPredictor(g, coeffs, intercepts, post_transform=post)
        gout = pred.forward_shared(pred.encode(x))
        oout, _ = po.linear(o, x, coeffs, intercepts, post_transform=post)
        assert same_rep(gout, oout)
        dec = g.fixedpoint_decode(40, g.reveal(gout)).to_host()
        z = x @ coeffs.T + intercepts
        if post == "softmax":
            e = np.exp(z - z.max(axis=1, keepdims=True))
            ref = e / e.sum(axis=1, keepdims=True)
        else:
            sg = 1.0 / (1.0 + np.exp(-z))
            ref = sg / sg.sum(axis=1, keepdims=True)
        np.testing.assert_almost_equal(dec, ref, decimal=2)
    finally:
        g.close()


@pytest.mark.parametrize("n_out", [1, 4])
def test_mlp_classifier_ring128_share_parity(n_out):
    """MLPClassifier post-transform (multilayer_perceptron_predictor.py:131-158): softmax over the classes, or
    [1 - sigmoid, sigmoid] for a single output; Ring128 fixed(24,40)."""
    rng = np.random.default_rng(12 + n_out)
    dims = [12, 8, n_out]
    ws = [rng.normal(0, dims[i] ** -0.5, (dims[i], dims[i + 1])) for i in range(2)]
    bs = [rng.normal(0, 0.1, dims[i + 1]) for i in range(2)]
    x = rng.uniform(0, 1, (10, dims[0]))
    g, o = mb.Session(0, SESSION_ID, ENTROPY_KEY), mo.Session(SESSION_ID, ENTROPY_KEY)
    try:
        pred = MLPPredictor(g, ws, bs, mb.RING128, 40, integral=24, post_transform="classifier")
        gout = pred.forward_shared(pred.encode(x))
        oout, _ = po.mlp(o, x, ws, bs, 128, 40, integral=24, post_transform="classifier")
        assert same_rep(gout, oout)
        dec = g.fixedpoint_decode(40, g.reveal(gout)).to_host()
        h = np.maximum(x @ ws[0] + bs[0], 0) @ ws[1] + bs[1]
        if n_out == 1:
            p = 1.0 / (1.0 + np.exp(-h))
            ref = np.concatenate([1 - p, p], axis=1)
        else:
            e = np.exp(h - h.max(axis=1, keepdims=True))
            ref = e / e.sum(axis=1, keepdims=True)
        np.testing.assert_almost_equal(dec, ref, decimal=2)
    finally:
        g.close()
