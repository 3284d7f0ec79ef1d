"""NeuralNetwork predictor (PyTorch / Keras exports) against the reference's own known answers
(tests/golden/nn_models.json from tests/golden/make_nn_golden.py; neural_network_predictor_test.py checks 2 decimals)."""
import json
import os

import numpy as np
import pytest

from _util import ENTROPY_KEY, SESSION_ID, same_rep

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "nn_models.json")))
X = np.array(GOLDEN["input"], dtype=np.float64)
MODELS = sorted(GOLDEN["models"])


def case(name):
    m = GOLDEN["models"][name]
    return [np.array(w) for w in m["weights"]], [np.array(b) for b in m["biases"]], m["activations"], np.array(m["expected"])


@pytest.mark.parametrize("name", MODELS)
def test_oracle_protocol_matches_reference_answers(name):
    from oracle import moose_oracle as mo
    from oracle import predictors_oracle as po
    ws, bs, acts, expected = case(name)
    _, got = po.neural_network(mo.Session(SESSION_ID, ENTROPY_KEY), X, ws, bs, acts)
    np.testing.assert_almost_equal(got, expected, decimal=2)


# Written after this round's GPU budget was spent: not yet executed on a GPU, so opt-in until it has been.
@pytest.mark.gpu
@pytest.mark.parametrize("name", MODELS)
def test_gpu_matches_oracle_bit_for_bit_and_reference_answers(name):
    import moose_b200 as mb
    from moose_b200.predictors import NeuralNetworkPredictor
    from oracle import moose_oracle as mo
    from oracle import predictors_oracle as po
    ws, bs, acts, expected = case(name)
    want_shares, want = po.neural_network(mo.Session(SESSION_ID, ENTROPY_KEY), X, ws, bs, acts)
    sess = mb.Session(0, SESSION_ID, ENTROPY_KEY)
    try:
        pred = NeuralNetworkPredictor(sess, ws, bs, acts)
        out = pred.forward_shared(pred.encode(X))
        assert same_rep(out, want_shares)
        got = sess.fixedpoint_decode(pred.frac, sess.reveal(out, 0)).to_host()
        assert np.array_equal(got, want)
        np.testing.assert_almost_equal(got, expected, decimal=2)
    finally:
        sess.close()
