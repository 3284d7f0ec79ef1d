"""GPU replay of the reference's own lowered computations through the per-op C ABI (moose_b200/executor.py):
tutorials/dotprod-compiled.moose (BASELINE configs[0]), its networked variant and the 19 045-op
moose/benches/rep_computation.moose.  Every operation's result is compared bit for bit with the oracle replay
run under the same session id and entropy key."""
import gzip
import os

import numpy as np
import pytest

import moose_b200 as mb
from moose_b200 import textual as tx
from moose_b200.executor import Executor, toposort
from oracle import moose_oracle as mo
from oracle import replay_oracle as ro
from _util import ENTROPY_KEY

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return tx.loads_json(gzip.open(os.path.join(GOLDEN, name + ".json.gz")).read().decode())


def same_value(g, o):
    if isinstance(g, mb.Tensor):
        if isinstance(o, mo.Ring):
            return g.shape == o.shape and np.array_equal(g.to_host(), o.data)
        if isinstance(o, mo.Bits):
            return g.shape == o.shape and np.array_equal(g.to_host(), o.packed)
        return np.array_equal(g.to_host(), np.asarray(o, dtype=np.float64))
    return g == o


@pytest.mark.parametrize("name,tag,expected", [("dotprod_compiled", "output_0", 32.0), ("dotprod_networked", "output_0", 32.0)])
def test_replay_dotprod(name, tag, expected):
    ops = load(name)
    sess = mb.Session(0, "replay", ENTROPY_KEY)
    try:
        ex = Executor(sess)
        outs = ex.run(ops)
        assert ex.executed == len(ops)
        got = outs[tag].to_host()
        assert got.shape == (1, 1) and abs(got[0, 0] - expected) < 1e-9
        ovals, _ = ro.replay(toposort(ops), mo.Session("replay", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16)
        for op in ops:
            assert same_value(ex.values[op.name], ovals[op.name]), op.name
    finally:
        sess.close()


def test_replay_rep_computation_19045_ops():
    ops = load("rep_computation")
    sess = mb.Session(0, "replay", ENTROPY_KEY)
    try:
        ex = Executor(sess)
        ex.run(ops)
        out = ex.saved["y_uri"].to_host()
        assert abs(out[0] - np.exp(2.0)) < 1e-5
        ovals, oouts = ro.replay(toposort(ops), mo.Session("replay", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16)
        assert np.array_equal(out, oouts["saved:y_uri"])
        mismatches = [op.name for op in ops if op.kind not in ("Save", "Output") and not same_value(ex.values[op.name], ovals[op.name])]
        assert not mismatches, mismatches[:5]
    finally:
        sess.close()


def test_replay_host_tail_graph():
    """Shape / indexing / plaintext-helper operators through the executor, op for op against the oracle replay."""
    from _util import HOST_TAIL_GRAPH
    ops = tx.parse_computation(HOST_TAIL_GRAPH)
    sess = mb.Session(0, "t", ENTROPY_KEY)
    try:
        ex = Executor(sess)
        outs = ex.run(ops)
        assert ex.executed == len(ops)
        ovals, oouts = ro.replay(toposort(ops), mo.Session("t", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16)
        for op in ops:
            assert same_value(ex.values[op.name], ovals[op.name]), op.name
        assert np.array_equal(outs["mean"].to_host(), oouts["mean"])
    finally:
        sess.close()


def test_replay_without_keeping_values():
    """keep_values=False: intermediates are released after their last use; outputs are unchanged."""
    ops = load("dotprod_networked")
    sess = mb.Session(0, "replay", ENTROPY_KEY)
    try:
        full = Executor(sess)
        want = full.run(ops)["output_0"].to_host()
        full.close()
        sess2 = mb.Session(0, "replay", ENTROPY_KEY)
        lean = Executor(sess2, keep_values=False)
        got = lean.run(ops)["output_0"].to_host()
        assert np.array_equal(got, want) and lean.executed == len(ops)
        with pytest.raises(KeyError):
            lean.values["op_20"]
        lean.close()
        sess2.close()
    finally:
        sess.close()


def test_replay_generated_mlp_graph():
    """The MLP graph emitted by moose_b200/lowering.py (ReLU via maximum: split, Kogge-Stone adder, dabit ring_inject,
    replicated mul, TruncPr) through the native executor: every operation bit for bit against the oracle replay, and
    the logits against a float MLP."""
    from moose_b200.lowering import Lowering, networking_pass, prune_pass
    from test_lowering import mlp_case
    net = networking_pass(prune_pass(Lowering(64).mlp_program(3, 23)))
    args, want = mlp_case(seed=5, dims=(9, 8, 6, 3), batch=7)
    sess = mb.Session(0, "t", ENTROPY_KEY)
    try:
        ex = Executor(sess)
        outs = ex.run(net, arguments={k: sess.from_host(mb.FLOAT64, v) for k, v in args.items()})
        assert ex.executed == len(net)
        got = outs["output_0"].to_host()
        assert np.abs(got - want).max() < 1e-5
        ovals, oouts = ro.replay(toposort(net), mo.Session("t", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16, args)
        assert np.array_equal(got, oouts["output_0"])
        bad = [op.name for op in net if op.kind not in ("Send", "Output") and not same_value(ex.values[op.name], ovals[op.name])]
        assert not bad, bad[:5]
    finally:
        sess.close()


# Written after round 1's GPU budget was spent: not yet executed on a GPU, so opt-in until it has been (tools/gpu_todo.sh).
def test_replay_generated_tree_regressor_graph():
    """The lowered + networked tree-ensemble regressor (moose_b200/lowering.py, about 10^4 host operations: compare =
    split + Kogge-Stone adder, mux = dabit ring_inject + replicated mul) through the native executor: every operation
    bit for bit against the oracle replay, the prediction against the reference's known answer for its fixture."""
    import json
    from moose_b200.lowering import Lowering, networking_pass, prune_pass
    from moose_b200.predictors import tree_ensemble_from_onnx_attributes
    golden = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tree_ensembles.json")))
    m = golden["models"]["xgboost_regressor"]
    kw = tree_ensemble_from_onnx_attributes(m["attributes"], m["n_features"], "regressor")
    net = networking_pass(prune_pass(Lowering(128).tree_regressor_program(kw["trees"], kw["n_features"], kw["base_score"], 40)))
    args = {"x": np.array(golden["input"], dtype=np.float64)}
    sess = mb.Session(0, "t", ENTROPY_KEY)
    try:
        ex = Executor(sess)
        outs = ex.run(net, arguments={k: sess.from_host(mb.FLOAT64, v) for k, v in args.items()})
        got = outs["output_0"].to_host()
        np.testing.assert_almost_equal(got, np.array(m["expected"]), decimal=2)
        ovals, oouts = ro.replay(toposort(net), mo.Session("t", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16, args)
        assert np.array_equal(got, oouts["output_0"])
        bad = [op.name for op in net if op.kind not in ("Send", "Output") and not same_value(ex.values[op.name], ovals[op.name])]
        assert not bad, bad[:5]
    finally:
        sess.close()
