"""Generates tests/golden/linear_models.json from the reference's fixtures and known answers (development container
only): LinearRegressor / LinearClassifier node attributes of pymoose/pymoose/predictors/fixtures/*.onnx and the expected
predictions of pymoose/pymoose/predictors/linear_predictor_test.py (_SK_REGRESSION_MODELS to 5 decimals for the input
[[1, 1, 1, 1], [-0.9, 1.3, 0.6, -0.4]], _SK_CLASSIFIER_MODELS to 2 decimals for [[-0.9, 1.3, 0.6, -0.4]])."""
import ast
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from moose_b200.onnx_lite import find_node  # noqa: E402

REF = "/root/reference/pymoose/pymoose/predictors"


def main():
    tree = ast.parse(open(os.path.join(REF, "linear_predictor_test.py")).read())
    golden = {"regression_input": [[1.0, 1.0, 1.0, 1.0], [-0.9, 1.3, 0.6, -0.4]], "classification_input": [[-0.9, 1.3, 0.6, -0.4]],
              "models": {}}
    for node in tree.body:
        if not (isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Name)):
            continue
        name = node.targets[0].id
        if name not in ("_SK_REGRESSION_MODELS", "_SK_CLASSIFIER_MODELS"):
            continue
        kind = "regressor" if "REGRESSION" in name else "classifier"
        for model, expected in ast.literal_eval(node.value):
            data = open(os.path.join(REF, "fixtures", model + ".onnx"), "rb").read()
            attrs, n_features = find_node(data, "LinearRegressor" if kind == "regressor" else "LinearClassifier")
            entry = {"kind": kind, "n_features": n_features, "expected": expected,
                     "coefficients": attrs["coefficients"]["floats"],
                     "intercepts": attrs["intercepts"]["floats"] if "intercepts" in attrs else None}
            if kind == "regressor":
                entry["targets"] = attrs["targets"]["i"] if "targets" in attrs else None
            else:
                labels = attrs.get("classlabels_ints", {}).get("ints") or attrs.get("classlabels_strings", {}).get("strings")
                entry["n_classes"] = len(labels)
                entry["post_transform"] = attrs["post_transform"]["s"]
            golden["models"][model] = entry
    path = os.path.join(ROOT, "tests", "golden", "linear_models.json")
    with open(path, "w") as f:
        json.dump(golden, f, separators=(",", ":"))
    print(len(golden["models"]), "models ->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
