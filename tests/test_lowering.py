"""The graph builder that restates the reference's lowering + networking passes for the replicated protocols
(moose_b200/lowering.py): same operator histogram and the same Send routes as the reference's own compiled graphs for
the same computation, and value-level correctness of the emitted MLP graph on the oracle's host kernels."""
import collections
import ctypes
import gzip
import os

import numpy as np
import pytest

from moose_b200 import textual as tx
from moose_b200.executor import toposort
from moose_b200.lowering import Lowering, networking_pass, prune_pass
from oracle import moose_oracle as mo
from oracle import replay_oracle as ro
from _util import ENTROPY_KEY

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return tx.loads_json(gzip.open(os.path.join(GOLDEN, name + ".json.gz")).read().decode())


def routes(ops):
    c = collections.Counter()
    for o in ops:
        if o.kind == "Send":
            c[(o.owners[0], o.attrs["receiver"].strip('"'), o.input_types[0])] += 1
    return c


def test_dot_graph_has_the_shape_of_the_reference_compilation():
    """tutorials/dotprod.moose compiled by the reference (tests/golden/dotprod_compiled, dotprod_networked) against
    the graph emitted here for the same computation: operator for operator the same counts, and the same
    (sender, receiver, type) multiset of Send operations."""
    ops = Lowering(128).fixed_dot_program([[1.0, 2.0, 3.0]], [[4.0], [5.0], [6.0]], 40)
    ref = load("dotprod_compiled")
    assert collections.Counter(o.kind for o in ops) == collections.Counter(o.kind for o in ref)
    assert collections.Counter((o.kind, o.owners[0]) for o in ops) == collections.Counter((o.kind, o.owners[0]) for o in ref)
    net, refn = networking_pass(ops), load("dotprod_networked")
    assert routes(net) == routes(refn) and len(net) == len(refn)
    vals, outs = ro.replay(toposort(net), mo.Session("t", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16)
    assert abs(outs["output_0"][0, 0] - 32.0) < 1e-9


def mlp_case(seed=0, dims=(6, 5, 4, 3), batch=4):
    rng = np.random.default_rng(seed)
    args = {"x": rng.uniform(0, 1, (batch, dims[0]))}
    for l in range(len(dims) - 1):
        args[f"w{l}"] = rng.normal(0, dims[l] ** -0.5, (dims[l], dims[l + 1]))
        args[f"b{l}"] = rng.normal(0, 0.1, (1, dims[l + 1]))
    h = args["x"]
    for l in range(len(dims) - 1):
        h = h @ args[f"w{l}"] + args[f"b{l}"]
        if l < len(dims) - 2:
            h = np.maximum(h, 0)
    return args, h


def test_mlp_graph_on_oracle_kernels():
    ops = Lowering(64).mlp_program(3, 23)
    net = networking_pass(prune_pass(ops))
    assert len(net) < len(ops) + 400 and all(o.placement_kind == "Host" for o in net)
    args, want = mlp_case()
    _, outs = ro.replay(toposort(net), mo.Session("t", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16, args)
    assert np.abs(outs["output_0"] - want).max() < 1e-5
    # the native reader accepts the emitted text and schedules it like the Python scheduler
    import moose_b200 as mb
    g = ctypes.c_void_p()
    assert mb.lib().mb_graph_parse(None, tx.dumps_text(net).encode(), ctypes.byref(g)) == 0
    n = mb.lib().mb_graph_num_ops(g)
    assert [mb.lib().mb_graph_op_name(g, i).decode() for i in range(n)] == [o.name for o in toposort(net)]
    mb.lib().mb_graph_free(g)


@pytest.mark.parametrize("name", ["xgboost_regressor", "random_forest_regressor", "gradient_boosting_regressor"])
def test_tree_regressor_graph_reproduces_reference_answers(name):
    """The lowered + networked tree-ensemble regressor (about 10^4 host operations) replayed on the oracle's kernels
    gives the known answers of the reference's tree_ensemble_test.py for its ONNX fixture."""
    import json
    import os
    from moose_b200.predictors import tree_ensemble_from_onnx_attributes
    golden = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "tree_ensembles.json")))
    m = golden["models"][name]
    kw = tree_ensemble_from_onnx_attributes(m["attributes"], m["n_features"], "regressor")
    ops = Lowering(128).tree_regressor_program(kw["trees"], kw["n_features"], kw["base_score"], 40)
    net = networking_pass(prune_pass(ops))
    assert all(o.placement_kind == "Host" for o in net)
    _, outs = ro.replay(toposort(net), mo.Session("t", ENTROPY_KEY), tx.parse_constant, tx.attr_hex16,
                        {"x": np.array(golden["input"], dtype=np.float64)})
    np.testing.assert_almost_equal(outs["output_0"], np.array(m["expected"]), decimal=2)
    if name == "xgboost_regressor":   # the native reader accepts the text and schedules it like the Python scheduler
        import moose_b200 as mb
        g = ctypes.c_void_p()
        assert mb.lib().mb_graph_parse(None, tx.dumps_text(net).encode(), ctypes.byref(g)) == 0
        n = mb.lib().mb_graph_num_ops(g)
        assert [mb.lib().mb_graph_op_name(g, i).decode() for i in range(n)] == [o.name for o in toposort(net)]
        mb.lib().mb_graph_free(g)
