"""Generates tests/golden/tree_ensembles.json from the reference's own fixtures and known answers (run in the
development container, where /root/reference exists; the tests read only the JSON):
  * tree-ensemble node attributes of pymoose/pymoose/predictors/fixtures/*.onnx, read with moose_b200/onnx_lite.py;
  * the expected predictions of pymoose/pymoose/predictors/tree_ensemble_test.py (lists _*_MODELS, parsed with ast)
    for the input [[0, 1, 1, 0], [1, 0, 1, 0], [0.2, 4, 2, 6]] (tree_ensemble_test.py:149-151), checked there to
    two decimals."""
import ast
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from moose_b200.onnx_lite import find_node  # noqa: E402

REF = "/root/reference/pymoose/pymoose/predictors"
KEEP = ["nodes_treeids", "nodes_truenodeids", "nodes_falsenodeids", "nodes_values", "nodes_featureids", "base_values",
        "target_treeids", "target_nodeids", "target_weights", "class_ids", "class_nodeids", "class_treeids", "class_weights",
        "classlabels_int64s", "classlabels_strings", "post_transform"]


def expected_tables():
    tree = ast.parse(open(os.path.join(REF, "tree_ensemble_test.py")).read())
    out = {}
    for node in tree.body:
        if isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Name):
            name = node.targets[0].id
            if name in ("_XGB_REGRESSOR_MODELS", "_SK_REGRESSOR_MODELS", "_XGB_CLASSIFIER_MODELS", "_SK_CLASSIFIER_MODELS"):
                kind = "regressor" if "REGRESSOR" in name else "classifier"
                for model, expected in ast.literal_eval(node.value):
                    out[model] = (kind, expected)
    return out


def main():
    golden = {"input": [[0, 1, 1, 0], [1, 0, 1, 0], [0.2, 4, 2, 6]], "models": {}}
    for model, (kind, expected) in sorted(expected_tables().items()):
        data = open(os.path.join(REF, "fixtures", model + ".onnx"), "rb").read()
        attrs, n_features = find_node(data, "TreeEnsembleRegressor" if kind == "regressor" else "TreeEnsembleClassifier")
        slim = {}
        for k in KEEP:
            if k in attrs:
                slim[k] = {f: v for f, v in attrs[k].items() if v not in ([], None)}
        golden["models"][model] = {"kind": kind, "n_features": n_features, "attributes": slim, "expected": expected}
    path = os.path.join(ROOT, "tests", "golden", "tree_ensembles.json")
    with open(path, "w") as f:
        json.dump(golden, f, separators=(",", ":"))
    print(len(golden["models"]), "models ->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
