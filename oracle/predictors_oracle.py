"""TEST INFRASTRUCTURE: the same predictor op sequences as moose_b200/predictors.py, on the oracle."""
import numpy as np

from . import moose_oracle as mo


def mlp_forward_shared(sess, x_fixed, w_fixed, b_fixed, frac, relu_via="maximum", integral=14, post_transform="none",
                       bias_via="share", activation="relu"):
    bits = x_fixed.bits
    x = mo.rep_share(sess, x_fixed, 0)
    n = len(w_fixed)
    for i in range(n):
        w = mo.rep_share(sess, w_fixed[i], 0)
        y = mo.rep_trunc_pr(sess, frac, mo.rep_dot(sess, x, w))
        b = b_fixed[i]
        # "share": AddOp::fixed_rep_kernel shares the mirrored bias from player 0 (fixedpoint/ops.rs:30-52)
        z = mo.rep_add_mir(y, mo.MirTensor([b, b, b])) if bias_via == "public" else mo.rep_add(y, mo.rep_share(sess, b, 0))
        if i < n - 1 and activation == "sigmoid":
            z = mo.rep_sigmoid(sess, integral, frac, z)
        elif i < n - 1 and activation == "relu":
            if relu_via == "maximum":
                zeros = mo.rep_share(sess, mo.fill(bits, 0, z.shape), 1)
                z = mo.rep_maximum(sess, [zeros, z])
            else:
                z = mo.rep_relu(sess, z)
        x = z
    if post_transform == "classifier":
        n_classes = b_fixed[-1].shape[-1]
        if n_classes == 1:
            pos = mo.rep_sigmoid(sess, integral, frac, x)
            if bias_via == "public":
                one = mo.fill(bits, 1 << frac, pos.shape)
                x = mo.rep_concat(1, [mo.rep_mir_sub(mo.MirTensor([one, one, one]), pos), pos])
            else:
                one = mo.rep_share(sess, mo.fixed_encode(bits, np.array(1.0), frac), 0)
                x = mo.rep_concat(1, [mo.rep_sub(one, pos), pos])
        else:
            x = mo.rep_softmax(sess, 1, n_classes, integral, frac, x)
    return x


def mlp(sess, x, weights, biases, bits=64, frac=23, relu_via="maximum", integral=14, post_transform="none", bias_via="share",
        activation="relu"):
    xf = mo.fixed_encode(bits, x, frac)
    wf = [mo.fixed_encode(bits, np.asarray(w, dtype=np.float64), frac) for w in weights]
    bf = [mo.fixed_encode(bits, np.asarray(b, dtype=np.float64).reshape(1, -1), frac) for b in biases]
    out = mlp_forward_shared(sess, xf, wf, bf, frac, relu_via, integral, post_transform, bias_via, activation)
    return out, mo.fixed_decode(mo.rep_reveal(out), frac)


def linear(sess, x, coeffs, intercepts, bits=128, frac=40, integral=24, post_transform="none"):
    w = np.concatenate([np.asarray(intercepts, dtype=np.float64).reshape(1, -1), np.asarray(coeffs, dtype=np.float64).T], axis=0)
    bias = mo.rep_share(sess, mo.fill(bits, 1 << frac, (x.shape[0], 1)), 1)
    xr = mo.rep_concat(1, [bias, mo.rep_share(sess, mo.fixed_encode(bits, x, frac), 0)])
    wr = mo.rep_share(sess, mo.fixed_encode(bits, w, frac), 0)
    y = mo.rep_trunc_pr(sess, frac, mo.rep_dot(sess, xr, wr))
    if post_transform == "sigmoid":
        y = mo.rep_sigmoid(sess, integral, frac, y)
    elif post_transform == "normalized_sigmoid":
        y = mo.rep_sigmoid(sess, integral, frac, y)
        y = mo.rep_div(sess, integral, frac, y, mo.rep_expand_dims([1], mo.rep_sum(1, y)))
    elif post_transform == "softmax":
        y = mo.rep_softmax(sess, 1, w.shape[1], integral, frac, y)
    return y, mo.fixed_decode(mo.rep_reveal(y), frac)


def tree_ensemble(sess, x, trees, n_features, base_score=0.0, learning_rate=1.0, n_classes=None, transform_output=False,
                  tree_class_map=None, bits=128, frac=40, integral=24):
    """pymoose/pymoose/predictors/tree_ensemble.py:33-62 (traversal), 83-114 (forest), 273-314 and 391-395 (post
    transforms) on the oracle, in graph-construction order; constants are rank-0 (predictor.py:19-23).  Host operands
    of replicated kernels are shared where the kernel shares them: x by alice in every index_axis
    (fixedpoint/ops.rs:1085-1104), mirrored thresholds / the mirrored one by player 0 (mirrored/ops.rs:159-180), leaf
    and base-score constants by carole (fixedpoint/ops.rs:2014-2051, 30-52)."""
    const = lambda v: mo.fixed_encode(bits, np.array(float(v)), frac)
    xf = mo.fixed_encode(bits, np.asarray(x, dtype=np.float64), frac)

    def traverse(tree, node, feats):
        if tree.left[node] == 0 and tree.right[node] == 0:
            return const(learning_rate * tree.weights[node])
        thr = mo.rep_share(sess, const(tree.split_conditions[node]), 0)
        selector = mo.rep_less(sess, feats[tree.split_indices[node]], thr)
        left = traverse(tree, tree.left[node], feats)
        right = traverse(tree, tree.right[node], feats)
        xs = left if isinstance(left, mo.RepTensor) else mo.rep_share(sess, left, 2)
        ys = right if isinstance(right, mo.RepTensor) else mo.rep_share(sess, right, 2)
        return mo.rep_mux(sess, mo.rep_ring_inject(sess, bits, 0, selector), xs, ys)

    scores = []
    for tree in trees:
        feats = [mo.rep_index_axis(1, i, mo.rep_share(sess, xf, 0)) for i in range(n_features)]
        scores.append(traverse(tree, 0, feats))
    base = lambda: mo.rep_share(sess, const(base_score), 2)
    if n_classes is None:
        total = mo.rep_add_n(scores)
        y = mo.rep_add(base(), total)
    elif n_classes == 2:
        logit = mo.rep_add(mo.rep_add_n(scores), base())
        pos = mo.rep_sigmoid(sess, integral, frac, logit) if transform_output else logit
        pos = mo.rep_expand_dims([1], pos)
        one = mo.rep_share(sess, const(1.0), 0)
        y = mo.rep_concat(1, [mo.rep_sub(one, pos), pos])
    else:
        per_class = [[] for _ in range(n_classes)]
        for tree_ix, class_ix in tree_class_map.items():
            per_class[class_ix].append(scores[tree_ix])
        logits = [mo.rep_add(mo.rep_add_n(v), base()) for v in per_class]
        y = mo.rep_concat(1, [mo.rep_expand_dims([1], v) for v in logits])
        if transform_output:
            y = mo.rep_softmax(sess, 1, n_classes, integral, frac, y)
    return y, mo.fixed_decode(mo.rep_reveal(y), frac)


def neural_network(sess, x, weights, biases, activations, bits=128, frac=40, integral=24):
    """pymoose/pymoose/predictors/neural_network_predictor.py:17-64 on the oracle: dot (mirrored weights shared by
    player 0), add (mirrored rank-1 bias shared by player 0), activation, on every layer."""
    n_classes = int(np.shape(biases[-1])[0])
    h = mo.rep_share(sess, mo.fixed_encode(bits, np.asarray(x, dtype=np.float64), frac), 0)
    for w, b, act in zip(weights, biases, activations):
        wr = mo.rep_share(sess, mo.fixed_encode(bits, np.asarray(w, dtype=np.float64), frac), 0)
        y = mo.rep_trunc_pr(sess, frac, mo.rep_dot(sess, h, wr))
        z = mo.rep_add(y, mo.rep_share(sess, mo.fixed_encode(bits, np.asarray(b, dtype=np.float64), frac), 0))
        if act == "sigmoid":
            z = mo.rep_sigmoid(sess, integral, frac, z)
        elif act == "relu":
            z = mo.rep_maximum(sess, [mo.rep_share(sess, mo.fill(bits, 0, z.shape), 1), z])
        elif act == "softmax":
            z = mo.rep_softmax(sess, 1, n_classes, integral, frac, z)
        h = z
    return h, mo.fixed_decode(mo.rep_reveal(h), frac)
