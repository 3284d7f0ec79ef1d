"""Host-side mirrors of the pymoose predictors on the path (op sequence only; all arithmetic is in
libmoose_b200).  They follow pymoose/pymoose/predictors/{multilayer_perceptron,linear}_predictor.py:

  MLP  (multilayer_perceptron_predictor.py:86-123): per layer  y = dot(x, w); z = y + b;
       hidden layers  z = maximum([zeros@bob, z])  (Activation.RELU, :98-111); logits out.
       The logical dot on fixed-point operands appends trunc_pr(f) (moose/src/logical/ops.rs:776-821).
  Linear (linear_predictor.py:33-63): bias = ones[B, 1] at bob; x' = concatenate([bias, x], axis=1) (both shared,
       in list order, then a share-wise concat: fixedpoint/ops.rs:1644-1667); y = dot(x', w) with the intercepts
       as the first row of w.  Post-transforms (linear_predictor.py:160-175, 294-301;
       multilayer_perceptron_predictor.py:131-158): sigmoid, softmax(axis=1, upmost_index=n_classes),
       normalized sigmoid (sigmoid / row sum), and MLPClassifier's [1 - p, p] for a single output.

Operands: x is a Host fixed-point tensor at alice (shared by party 0 inside the first dot,
fixedpoint/ops.rs:788-812); weights and biases are mirrored fixed-point constants (weights are shared
from mirrored player 0, mirrored/ops.rs:154-180; so are the biases inside AddOp, fixedpoint/ops.rs:30-52);
the ReLU zeros are a Host tensor at bob (shared by party 1).
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np

from . import FLOAT64, RING64, RING128, RepTensor, Session, Tensor


class MLPPredictor:
    def __init__(self, sess: Session, weights: Sequence[np.ndarray], biases: Sequence[np.ndarray],
                 ring: int = RING64, frac: int = 23, relu_via: str = "maximum", integral: int = 14,
                 post_transform: str = "none", bias_via: str = "share", activation: str = "relu"):
        """post_transform: "none" (MLPRegressor) or "classifier" (MLPClassifier: softmax over the classes, or
        [1 - sigmoid, sigmoid] for one output; needs 2 * (integral + frac) <= ring bits, division.rs:33-38).
        bias_via: "share" is what the logical graph does -- AddOp's replicated kernel on abstract fixed tensors shares a
        mirrored operand from player 0 first (fixedpoint/ops.rs:30-52, mirrored/ops.rs:159-180) -- and "public" adds
        it to share 0 only (the concrete rep + mirrored kernel, arith.rs:81-121): same revealed value, one PRF
        expansion and two additions fewer.  "share", the reference's sequence, is the default and what bench.py measures."""
        self.sess, self.ring, self.frac, self.relu_via = sess, ring, frac, relu_via
        self.bias_via = bias_via
        if activation not in ("relu", "sigmoid", "identity"):   # Activation enum, multilayer_perceptron_predictor.py:69-78
            raise ValueError("Invalid or unsupported activation function")
        self.activation = activation
        self.integral, self.post = integral, post_transform
        enc = lambda a: sess.fixedpoint_encode(ring, frac, sess.from_host(FLOAT64, np.asarray(a, dtype=np.float64)))
        self.w: List[Tensor] = [enc(w) for w in weights]                      # [in, out] each
        self.b: List[Tensor] = [enc(np.asarray(b).reshape(1, -1)) for b in biases]

    def encode(self, x: np.ndarray) -> Tensor:
        return self.sess.fixedpoint_encode(self.ring, self.frac, self.sess.from_host(FLOAT64, x))

    def forward_shared(self, x_fixed: Tensor) -> RepTensor:
        s = self.sess
        x = s.share(x_fixed, 0)
        n = len(self.w)
        for i in range(n):
            w = s.share(self.w[i], 0)
            y = s.fixed_dot(self.frac, x, w)
            z = s.rep_add_public(y, self.b[i]) if self.bias_via == "public" else s.rep_add(y, s.share(self.b[i], 0))
            if i < n - 1 and self.activation == "sigmoid":
                z = s.sigmoid(self.integral, self.frac, z)
            elif i < n - 1 and self.activation == "relu":
                if self.relu_via == "maximum":
                    zeros = s.share(s.fill(self.ring, 0, z.shape), 1)
                    z = s.maximum([zeros, z])
                else:
                    z = s.relu(z)
            x = z
        if self.post == "classifier":
            n_classes = self.b[-1].shape[-1]
            if n_classes == 1:
                pos = s.sigmoid(self.integral, self.frac, x)
                if self.bias_via == "public":
                    one = s.fill(self.ring, 1 << self.frac, pos.shape)
                    x = s.rep_concat(1, [s.rep_public_sub(one, pos), pos])
                else:   # the mirrored rank-0 one is shared by player 0 inside SubOp (fixedpoint/ops.rs, SubOp::fixed_rep_kernel)
                    one = s.share(s.fixedpoint_encode(self.ring, self.frac, s.from_host(FLOAT64, np.array(1.0))), 0)
                    x = s.rep_concat(1, [s.rep_sub(one, pos), pos])
            else:
                x = s.softmax(1, n_classes, self.integral, self.frac, x)
        return x

    def __call__(self, x: np.ndarray) -> np.ndarray:
        out = self.forward_shared(self.encode(x))
        return self.sess.fixedpoint_decode(self.frac, self.sess.reveal(out, 0)).to_host()


class LinearPredictor:
    """LinearRegressor (post_transform "none") / LinearClassifier with "sigmoid" (two classes), "normalized_sigmoid"
    (more classes) or "softmax" (linear_predictor.py:160-175)."""

    def __init__(self, sess: Session, coeffs: np.ndarray, intercepts: np.ndarray, ring: int = RING128, frac: int = 40,
                 integral: int = 24, post_transform: str = "none"):
        self.sess, self.ring, self.frac, self.integral, self.post = sess, ring, frac, integral, post_transform
        w = np.concatenate([np.asarray(intercepts, dtype=np.float64).reshape(1, -1), np.asarray(coeffs, dtype=np.float64).T], axis=0)
        self.n_classes = w.shape[1]
        self.w = sess.fixedpoint_encode(ring, frac, sess.from_host(FLOAT64, w))  # [1 + features, classes]

    def encode(self, x: np.ndarray) -> Tensor:
        return self.sess.fixedpoint_encode(self.ring, self.frac, self.sess.from_host(FLOAT64, x))

    def forward_shared(self, x_fixed: Tensor) -> RepTensor:
        """x_fixed: [B, features] fixed-point host tensor at alice (no bias column)."""
        s = self.sess
        bias = s.share(s.fill(self.ring, 1 << self.frac, (x_fixed.shape[0], 1)), 1)  # bias_trick at bob
        x = s.share(x_fixed, 0, deferred=True)   # the [B, features] sharing is expanded inside the dot, never stored
        w = s.share(self.w, 0)
        y = s.fixed_dot_concat(self.frac, [bias, x], w)  # = trunc_pr(dot(concatenate([bias, x], axis=1), w))
        if self.post == "sigmoid":
            y = s.sigmoid(self.integral, self.frac, y)
        elif self.post == "normalized_sigmoid":
            y = s.sigmoid(self.integral, self.frac, y)
            y = s.rep_div(self.integral, self.frac, y, s.rep_expand_dims(1, s.rep_sum(y, 1)))
        elif self.post == "softmax":
            y = s.softmax(1, self.n_classes, self.integral, self.frac, y)
        elif self.post != "none":
            raise ValueError(f"unknown post_transform {self.post!r}")
        return y

    def __call__(self, x: np.ndarray) -> np.ndarray:
        out = self.forward_shared(self.encode(np.asarray(x, dtype=np.float64)))
        return self.sess.fixedpoint_decode(self.frac, self.sess.reveal(out, 0)).to_host()


# ---- tree ensembles (pymoose/pymoose/predictors/tree_ensemble.py) -------------------------------------------------------
class DecisionTree:
    """DecisionTreeRegressor (tree_ensemble.py:8-62): node arrays in ONNX order; a node with left == right == 0 is a
    leaf whose value is weights[node]."""

    def __init__(self, weights, children, split_conditions, split_indices):
        self.weights = dict(weights)
        self.left, self.right = list(children[0]), list(children[1])
        self.split_conditions, self.split_indices = list(split_conditions), list(split_indices)

    def is_leaf(self, node: int) -> bool:
        return self.left[node] == 0 and self.right[node] == 0

    def evaluate_float(self, row, node: int = 0, scale: float = 1.0) -> float:
        """Plaintext walk (tests): x[split] < threshold goes left (tree_ensemble.py:41-55)."""
        while not self.is_leaf(node):
            node = self.left[node] if row[self.split_indices[node]] < self.split_conditions[node] else self.right[node]
        return scale * self.weights[node]


def tree_ensemble_from_onnx_attributes(attrs: dict, n_features: int, kind: str):
    """The reference's from_onnx (tree_ensemble.py:139-270, 329-389, 398-466) on the attribute dictionary of the
    TreeEnsembleRegressor / TreeEnsembleClassifier node ({name: {"ints" | "floats" | "strings" | "s"}},
    moose_b200/onnx_lite.py).  Returns the keyword arguments of TreeEnsemblePredictor."""
    ints = lambda n: list(attrs[n]["ints"])
    floats = lambda n: list(attrs[n]["floats"])
    nodes_treeids, left, right = ints("nodes_treeids"), ints("nodes_truenodeids"), ints("nodes_falsenodeids")
    conds, feats = floats("nodes_values"), ints("nodes_featureids")
    n_trees = len(set(nodes_treeids))
    if len(set(feats)) > n_features or max(feats) > n_features:
        raise ValueError("split indices exceed the input's feature count")
    base = attrs.get("base_values", {}).get("floats") or [0.0]
    out = {"n_features": n_features, "base_score": float(base[0]), "learning_rate": 1.0}
    if kind == "regressor":
        final_nodes, w_tree, w_node, w_val = nodes_treeids, ints("target_treeids"), ints("target_nodeids"), floats("target_weights")
        class_ids = None
    else:
        class_ids, w_node = ints("class_ids"), ints("class_nodeids")
        class_treeids, w_val = ints("class_treeids"), floats("class_weights")
        labels = attrs.get("classlabels_int64s", {}).get("ints") or attrs.get("classlabels_strings", {}).get("strings")
        n_classes = len(labels)
        post = attrs["post_transform"]["s"]
        if post == "NONE" and n_classes > 2:
            # sklearn stores one tree with an n_classes vector per node: one copy of every tree per class
            w_tree = [c + t * n_classes for t, c in zip(class_treeids, class_ids)]
            n_trees = len(set(w_tree))
            assert nodes_treeids == sorted(nodes_treeids)
            sub = [[t for t in nodes_treeids if t == i] for i in sorted(set(nodes_treeids))]
            final_nodes = [n_classes * i + j for j in range(n_classes) for i, x in enumerate(sub) for _ in x]
        else:
            w_tree, final_nodes = class_treeids, nodes_treeids
        out.update(n_classes=n_classes, transform_output=post != "NONE")
    args = [{"weights": {}, "children": [[], []], "split_indices": [], "split_conditions": []} for _ in range(n_trees)]
    for i, t in enumerate(final_nodes):
        args[t]["children"][0].append(left[i % len(left)])
        args[t]["children"][1].append(right[i % len(right)])
        args[t]["split_indices"].append(feats[i % len(feats)])
        args[t]["split_conditions"].append(conds[i % len(conds)])
    for i, w in enumerate(w_val):
        args[w_tree[i]]["weights"][w_node[i]] = w
    out["trees"] = [DecisionTree(**a) for a in args]
    if class_ids is not None:
        out["tree_class_map"] = {t: c for t, c in zip(w_tree, class_ids)}
    return out


class TreeEnsemblePredictor:
    """TreeEnsembleRegressor (n_classes None) / TreeEnsembleClassifier (tree_ensemble.py:65-114, 117-314, 317-395).
    Per tree and inner node, in the order the reference's graph is built: selector = less(x[:, split], threshold) with
    the mirrored threshold shared by player 0 (fixedpoint/ops.rs:1815-1845, mirrored/ops.rs:154-180), then the left
    and the right subtree, then mux(selector, left, right) = ring_inject(selector) * (left - right) + right with leaf
    constants shared by carole (fixedpoint/ops.rs:2014-2051, replicated/control_flow.rs:27-44).  Constants are rank-0
    tensors (predictor.py:19-23) and broadcast against the [batch] feature columns.  A host input is shared by every
    index_axis that reads it (fixedpoint/ops.rs:1085-1104), as in the reference's own test flow; a replicated input
    (the AES flow) is used as is."""

    def __init__(self, sess: Session, trees: Sequence[DecisionTree], n_features: int, base_score: float = 0.0,
                 learning_rate: float = 1.0, n_classes=None, transform_output: bool = False, tree_class_map=None,
                 ring: int = RING128, frac: int = 40, integral: int = 24):
        self.sess, self.trees, self.n_features = sess, list(trees), n_features
        self.base_score, self.learning_rate = base_score, learning_rate
        self.n_classes, self.transform_output, self.tree_class_map = n_classes, transform_output, tree_class_map
        self.ring, self.frac, self.integral = ring, frac, integral

    @classmethod
    def from_onnx(cls, sess: Session, model_bytes: bytes, kind: str, **kw):
        from .onnx_lite import find_node
        attrs, n_features = find_node(model_bytes, "TreeEnsembleRegressor" if kind == "regressor" else "TreeEnsembleClassifier")
        if attrs is None:
            raise ValueError("Incompatible ONNX graph provided: no tree ensemble operator of that kind")
        return cls(sess, **tree_ensemble_from_onnx_attributes(attrs, n_features, kind), **kw)

    def encode(self, x: np.ndarray) -> Tensor:
        return self.sess.fixedpoint_encode(self.ring, self.frac, self.sess.from_host(FLOAT64, np.asarray(x, dtype=np.float64)))

    def _const(self, value: float) -> Tensor:
        return self.sess.fixedpoint_encode(self.ring, self.frac, self.sess.from_host(FLOAT64, np.array(float(value))))

    def _traverse(self, tree: DecisionTree, node: int, feats):
        """Replicated tensor for an inner node, host constant (to be shared by carole at its use) for a leaf."""
        s = self.sess
        if tree.is_leaf(node):
            return self._const(self.learning_rate * tree.weights[node])
        thr = s.share(self._const(tree.split_conditions[node]), 0)
        selector = s.rep_less(feats[tree.split_indices[node]], thr)
        left = self._traverse(tree, tree.left[node], feats)
        right = self._traverse(tree, tree.right[node], feats)
        x = left if isinstance(left, RepTensor) else s.share(left, 2)
        y = right if isinstance(right, RepTensor) else s.share(right, 2)
        return s.mux(s.rep_ring_inject(self.ring, 0, selector), x, y)

    def tree_scores(self, x) -> List[RepTensor]:
        s, scores = self.sess, []
        for tree in self.trees:
            feats = [s.rep_index_axis(1, i, x if isinstance(x, RepTensor) else s.share(x, 0)) for i in range(self.n_features)]
            score = self._traverse(tree, 0, feats)
            if not isinstance(score, RepTensor):
                raise NotImplementedError("degenerate tree (a single leaf)")
            scores.append(score)
        return scores

    def forward_shared(self, x) -> RepTensor:
        s = self.sess
        scores = self.tree_scores(x)
        base = lambda: s.share(self._const(self.base_score), 2)
        if self.n_classes is None:                                   # regressor: add(base_score, add_n(scores))
            total = s.rep_add_n(scores)
            return s.rep_add(base(), total)
        if self.n_classes == 2:                                      # _maybe_sigmoid
            logit = s.rep_add(s.rep_add_n(scores), base())
            pos = s.sigmoid(self.integral, self.frac, logit) if self.transform_output else logit
            pos = s.rep_expand_dims(1, pos)
            one = s.share(self._const(1.0), 0)                       # mirrored constant, shared inside SubOp
            return s.rep_concat(1, [s.rep_sub(one, pos), pos])
        per_class = [[] for _ in range(self.n_classes)]              # _ovr_logit
        for tree_ix, class_ix in self.tree_class_map.items():
            per_class[class_ix].append(scores[tree_ix])
        logits = [s.rep_add(s.rep_add_n(v), base()) for v in per_class]
        logit = s.rep_concat(1, [s.rep_expand_dims(1, v) for v in logits])
        return s.softmax(1, self.n_classes, self.integral, self.frac, logit) if self.transform_output else logit

    def __call__(self, x: np.ndarray) -> np.ndarray:
        out = self.forward_shared(self.encode(x))
        return self.sess.fixedpoint_decode(self.frac, self.sess.reveal(out, 0)).to_host()


# ---- generic dense networks (pymoose/pymoose/predictors/neural_network_predictor.py) ------------------------------------
def neural_network_from_onnx(model_bytes: bytes):
    """NeuralNetwork.from_onnx (neural_network_predictor.py:66-153): (weights [in, out] per layer, biases, activations)
    of a PyTorch (Gemm) or TF-Keras (MatMul + BiasAdd) export."""
    from . import onnx_lite as ol
    ops = [op for _, op, _ in ol.nodes(model_bytes)]
    acts = []
    for i, op in enumerate(ops):
        if op in ("Sigmoid", "Softmax", "Relu"):
            acts.append(op.lower())
        if i > 0 and op == "Gemm" and ops[i - 1] == "Gemm":
            acts.append("identity")
        if i > 2 and ops[i - 3:i + 1] == ["MatMul", "Add", "MatMul", "Add"]:
            acts.append("identity")
    ini = ol.initializers(model_bytes)
    weights, biases = [], []
    for name, dims, dtype, vals in ini:
        if "weight" in name or "MatMul" in name:
            if dtype != 1:
                raise ValueError("Neural Network Weights must be of type FLOATS, found other.")
            weights.append(np.asarray(vals[:dims[0] * dims[1]], dtype=np.float64).reshape(dims[0], dims[1]).T)
    for name, dims, dtype, vals in ini:
        if "bias" in name or "BiasAdd" in name:
            if dtype != 1:
                raise ValueError("Neural network biases must be of type FLOATS, found other.")
            biases.append(np.asarray(vals[:dims[0]], dtype=np.float64))
    if "tf" in ol.producer_name(model_bytes):      # Keras stores the layers last to first, weights transposed
        weights, biases = [w.T for w in weights[::-1]], biases[::-1]
    if ol.input_features(model_bytes) != weights[0].shape[0]:
        raise ValueError("the input shape of the ONNX file does not match the first layer's weights")
    return weights, biases, acts


class NeuralNetworkPredictor:
    """NeuralNetwork (neural_network_predictor.py:17-64): per layer dot with the mirrored weights (shared by player 0
    inside DotOp), add of the mirrored bias (shared by player 0 inside AddOp, fixedpoint/ops.rs:30-52), then that
    layer's activation -- sigmoid, relu = maximum([zeros@bob, z]), softmax(axis 1, upmost_index = n_classes) or none --
    on EVERY layer, the last included."""

    def __init__(self, sess: Session, weights, biases, activations, ring: int = RING128, frac: int = 40, integral: int = 24):
        self.sess, self.ring, self.frac, self.integral = sess, ring, frac, integral
        enc = lambda a: sess.fixedpoint_encode(ring, frac, sess.from_host(FLOAT64, np.asarray(a, dtype=np.float64)))
        self.w = [enc(w) for w in weights]
        self.b = [enc(np.asarray(b)) for b in biases]              # rank 1, broadcast over the batch
        self.activations = list(activations)
        self.n_classes = int(np.shape(biases[-1])[0])

    def encode(self, x: np.ndarray) -> Tensor:
        return self.sess.fixedpoint_encode(self.ring, self.frac, self.sess.from_host(FLOAT64, np.asarray(x, dtype=np.float64)))

    def forward_shared(self, x_fixed: Tensor) -> RepTensor:
        s = self.sess
        x = s.share(x_fixed, 0)
        for i in range(len(self.w)):
            y = s.fixed_dot(self.frac, x, s.share(self.w[i], 0))
            z = s.rep_add(y, s.share(self.b[i], 0))
            act = self.activations[i]
            if act == "sigmoid":
                z = s.sigmoid(self.integral, self.frac, z)
            elif act == "relu":
                z = s.maximum([s.share(s.fill(self.ring, 0, z.shape), 1), z])
            elif act == "softmax":
                z = s.softmax(1, self.n_classes, self.integral, self.frac, z)
            elif act != "identity":
                raise ValueError("Invalid or unsupported activation function")
            x = z
        return x

    def __call__(self, x: np.ndarray) -> np.ndarray:
        out = self.forward_shared(self.encode(x))
        return self.sess.fixedpoint_decode(self.frac, self.sess.reveal(out, 0)).to_host()
