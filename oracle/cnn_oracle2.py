"""A second, independently written CPU forward of the reference's networks -- TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED against Keras itself (Keras 1.2.0 / Theano 0.8.2 are not installable here); what
this file adds is a cross-check of oracle/cnn_oracle.py's reading of their semantics, written from
the layer definitions alone and sharing no code or library routine with it:

  * Convolution2D(border_mode='same', dim_ordering='th') on the Theano backend is
    theano.tensor.nnet.conv2d(input, filters, border_mode='half', filter_flip=True)
    (keras/backend/theano_backend.py, conv2d): a TRUE convolution
        out[n, o, x, y] = b[o] + sum_c sum_{i,j} in[n, c, x - i + p, y - j + p] * W[o, c, i, j],  p = k // 2
    which is exactly scipy.signal.convolve2d(in[n, c], W[o, c], mode='same') summed over c.
  * Flatten of (N, 1, S, S) is C order: index x*S + y.  Bias adds a vector (nn_util.py:118-133).
    softmax over the S*S entries (policy.py:134-136).  Dense is x.W + b with W (in, out) (value.py:27-28).
  * BatchNormalization(mode=0, axis=-1) at test time (keras/layers/normalization.py, call():
    broadcast_shape has input_shape[axis] at `axis` and 1 elsewhere; K.batch_normalization(x,
    running_mean, running_std, beta, gamma, epsilon) = (x - mean) / sqrt(std + eps) * gamma + beta,
    `running_std` holding the variance): with axis=-1 and 'th' ordering the four vectors have
    length S and vary along y (policy.py:226-228 builds it with the defaults).

Slow (pure numpy/scipy loops); use on small batches.
"""
import numpy as np
from scipy.signal import convolve2d


def conv_same(x, W, b):
    """x (N, C, S, S), W (O, C, k, k), b (O,) -> (N, O, S, S): Theano conv2d, border_mode 'half'."""
    x = np.asarray(x, np.float64)
    W = np.asarray(W, np.float64)
    N, C, S, _ = x.shape
    O = W.shape[0]
    out = np.empty((N, O, S, S), np.float64)
    for n in range(N):
        for o in range(O):
            acc = np.full((S, S), float(b[o]))
            for c in range(C):
                acc += convolve2d(x[n, c], W[o, c], mode="same")
            out[n, o] = acc
    return out


def softmax(z):
    z = z - z.max(axis=1, keepdims=True)
    e = np.exp(z)
    return e / e.sum(axis=1, keepdims=True)


def forward(weights, x, specs, head="policy"):
    """Sequential stack of policy.py:94-138 / value.py:14-28.  float64 arithmetic, float32 result."""
    h = np.asarray(x, np.float64)
    i = 0
    for _nb, _k in specs:
        h = np.maximum(conv_same(h, weights[i], weights[i + 1]), 0.0)
        i += 2
    h = conv_same(h, weights[i], weights[i + 1])
    i += 2
    flat = h.reshape(h.shape[0], -1)
    if head == "policy":
        return softmax(flat + np.asarray(weights[i], np.float64)).astype(np.float32)
    d1 = flat.dot(np.asarray(weights[i], np.float64)) + weights[i + 1]
    d2 = d1.dot(np.asarray(weights[i + 2], np.float64)) + weights[i + 3]
    return np.tanh(d2).reshape(-1).astype(np.float32)


def batchnorm_last_axis(x, gamma, beta, mean, var, eps=1e-3):
    shape = [1] * x.ndim
    shape[-1] = x.shape[-1]
    g, b, m, v = (np.asarray(a, np.float64).reshape(shape) for a in (gamma, beta, mean, var))
    return (x - m) / np.sqrt(v + eps) * g + b


def forward_resnet(weights, x, specs, unit_end):
    """ResnetPolicy.create_network (policy.py:141-271) at test time; weights in get_weights() order."""
    h = np.asarray(x, np.float64)
    path = conv_same(h, weights[0], weights[1])
    block_input = path
    i = 2
    for c in range(1, len(specs)):
        path = batchnorm_last_axis(path, *weights[i:i + 4])
        path = np.maximum(path, 0.0)
        path = conv_same(path, weights[i + 4], weights[i + 5])
        i += 6
        if unit_end[c]:
            path = block_input + path          # merge(mode='sum'), policy.py:240-245
            block_input = path
    path = np.maximum(path, 0.0)
    path = conv_same(path, weights[i], weights[i + 1])
    flat = path.reshape(path.shape[0], -1) + np.asarray(weights[i + 2], np.float64)
    return softmax(flat).astype(np.float32)
