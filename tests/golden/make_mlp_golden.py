"""Generates tests/golden/mlp_models.json.gz from the reference's fixtures and known answers (development container
only): the "coefficient*" / "intercepts*" initializers and the activation of the sklearn MLP ONNX fixtures
(multilayer_perceptron_predictor.py:25-80) and the expected predictions of
pymoose/pymoose/predictors/multilayer_perceptron_predictor_test.py (regressors to 1 decimal, classifiers to 2)."""
import ast
import base64
import gzip
import json
import os
import struct
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from moose_b200 import onnx_lite as ol  # noqa: E402

REF = "/root/reference/pymoose/pymoose/predictors"
ROW = [-0.9, 1.3, 0.6, -0.4, 0.8, 0.2, -0.1, -0.9, 0.5, 0.7, -0.6, 0.3, 2.3, 1.5, 1.9, -1.0, 1.0, 0.6, 1.4, -1.2]


def f32(values):
    return base64.b64encode(struct.pack(f"<{len(values)}f", *values)).decode()


def main():
    tree = ast.parse(open(os.path.join(REF, "multilayer_perceptron_predictor_test.py")).read())
    golden = {"regression_input": [[1.0] * 20, ROW], "classification_input": [ROW], "models": {}}
    for node in tree.body:
        if not (isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Name)):
            continue
        name = node.targets[0].id
        if name not in ("_SK_REGRESSION_MODELS", "_SK_CLASSIFIER_MODELS"):
            continue
        kind = "regressor" if "REGRESSION" in name else "classifier"
        for model, expected in ast.literal_eval(node.value):
            data = open(os.path.join(REF, "fixtures", model + ".onnx"), "rb").read()
            ini = ol.initializers(data)
            assert all(t == 1 for n, _, t, _ in ini if "coefficient" in n or "intercepts" in n)
            weights = [{"dims": d, "f32": f32(v)} for n, d, _, v in ini if "coefficient" in n]
            biases = [{"f32": f32(v)} for n, _, _, v in ini if "intercepts" in n]
            act = [n for n, _, outs in ol.nodes(data) if outs and outs[0] == "next_activations"]
            activation = {"Sigmoid": "sigmoid", "Relu": "relu"}.get(act[-1] if act else None, "identity")
            golden["models"][model] = {"kind": kind, "n_features": ol.input_features(data), "weights": weights, "biases": biases,
                                       "activation": activation, "expected": expected}
    path = os.path.join(ROOT, "tests", "golden", "mlp_models.json.gz")
    with gzip.GzipFile(path, "wb", mtime=0) as f:
        f.write(json.dumps(golden, separators=(",", ":")).encode())
    print(len(golden["models"]), "models ->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
