"""Generates tests/golden/nn_models.json from the reference's PyTorch / Keras ONNX fixtures and the known answers of
pymoose/pymoose/predictors/neural_network_predictor_test.py (development container only).  The layers are stored as
NeuralNetwork.from_onnx yields them (moose_b200.predictors.neural_network_from_onnx restates that conversion)."""
import ast
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from moose_b200.predictors import neural_network_from_onnx  # noqa: E402

REF = "/root/reference/pymoose/pymoose/predictors"


def main():
    src = open(os.path.join(REF, "neural_network_predictor_test.py")).read()
    tree = ast.parse(src)
    golden = {"models": {}}
    for node in ast.walk(tree):
        if isinstance(node, ast.Assign) and isinstance(node.targets[0], ast.Name):
            if node.targets[0].id == "_MODELS":
                models = ast.literal_eval(node.value)
            elif node.targets[0].id == "input_x":      # np.array([[...]], dtype=np.float64)
                golden["input"] = ast.literal_eval(node.value.args[0])
    for model, expected in models:
        ws, bs, acts = neural_network_from_onnx(open(os.path.join(REF, "fixtures", model + ".onnx"), "rb").read())
        golden["models"][model] = {"weights": [w.tolist() for w in ws], "biases": [b.tolist() for b in bs], "activations": acts,
                                   "expected": expected}
    path = os.path.join(ROOT, "tests", "golden", "nn_models.json")
    with open(path, "w") as f:
        json.dump(golden, f, separators=(",", ":"))
    print(len(golden["models"]), "models ->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
