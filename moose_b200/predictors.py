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
from mirrored player 0, mirrored/ops.rs:154-180; biases are added as public values, arith.rs:81-121);
the ReLU zeros are a Host tensor at bob (shared by party 1).
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np

from . import FLOAT64, RING64, RING128, RepTensor, Session, Tensor


class MLPPredictor:
    def __init__(self, sess: Session, weights: Sequence[np.ndarray], biases: Sequence[np.ndarray],
                 ring: int = RING64, frac: int = 23, relu_via: str = "maximum", integral: int = 14,
                 post_transform: str = "none"):
        """post_transform: "none" (MLPRegressor) or "classifier" (MLPClassifier: softmax over the classes, or
        [1 - sigmoid, sigmoid] for one output; needs 2 * (integral + frac) <= ring bits, division.rs:33-38)."""
        self.sess, self.ring, self.frac, self.relu_via = sess, ring, frac, relu_via
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
            y = s.trunc_pr(self.frac, s.rep_dot(x, w))
            z = s.rep_add_public(y, self.b[i])
            if i < n - 1:
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
                one = s.fill(self.ring, 1 << self.frac, pos.shape)
                x = s.rep_concat(1, [s.rep_public_sub(one, pos), pos])
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
        x = s.share(x_fixed, 0)
        w = s.share(self.w, 0)
        y = s.trunc_pr(self.frac, s.rep_dot_concat([bias, x], w))  # = dot(concatenate([bias, x], axis=1), w)
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
