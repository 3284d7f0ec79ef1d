"""TEST INFRASTRUCTURE: the same predictor op sequences as moose_b200/predictors.py, on the oracle."""
import numpy as np

from . import moose_oracle as mo


def mlp_forward_shared(sess, x_fixed, w_fixed, b_fixed, frac, relu_via="maximum", integral=14, post_transform="none"):
    bits = x_fixed.bits
    x = mo.rep_share(sess, x_fixed, 0)
    n = len(w_fixed)
    for i in range(n):
        w = mo.rep_share(sess, w_fixed[i], 0)
        y = mo.rep_trunc_pr(sess, frac, mo.rep_dot(sess, x, w))
        b = b_fixed[i]
        z = mo.rep_add_mir(y, mo.MirTensor([b, b, b]))
        if i < n - 1:
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
            one = mo.fill(bits, 1 << frac, pos.shape)
            x = mo.rep_concat(1, [mo.rep_mir_sub(mo.MirTensor([one, one, one]), pos), pos])
        else:
            x = mo.rep_softmax(sess, 1, n_classes, integral, frac, x)
    return x


def mlp(sess, x, weights, biases, bits=64, frac=23, relu_via="maximum", integral=14, post_transform="none"):
    xf = mo.fixed_encode(bits, x, frac)
    wf = [mo.fixed_encode(bits, np.asarray(w, dtype=np.float64), frac) for w in weights]
    bf = [mo.fixed_encode(bits, np.asarray(b, dtype=np.float64).reshape(1, -1), frac) for b in biases]
    out = mlp_forward_shared(sess, xf, wf, bf, frac, relu_via, integral, post_transform)
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
