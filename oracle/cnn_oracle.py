"""CPU restatement of the reference's network forward -- TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the arithmetic of `CNNPolicy.forward` lives in third-party Keras==1.2.0 on
Theano==0.8.2 (requirements.txt:3,10 of the reference), neither vendored nor installable
here, and the reference's own tests check shapes only (tests/test_policy.py:17-30).  This file
restates the published semantics of the layers the reference composes, in torch fp32 on CPU:

  policy.py:94-138   Convolution2D(k, k, 'same', relu) x layers, Convolution2D(1, 1), Flatten,
                     Bias, softmax
  value.py:14-28     the same stack, then Flatten, Dense(256) (linear), Dense(1, tanh)
  nn_util.py:118-133 Bias: adds a learned vector of shape input_shape[1:] (zeros initially)

Keras-1.2.0/Theano rules used: kernels are stored (out, in, kh, kw) and the
Theano backend performs a TRUE convolution (filter_flip=True), i.e. cross-correlation with the
kernel rotated 180 degrees; border_mode='same' with odd k pads k//2 zeros each side;
init='uniform' draws U(-0.05, 0.05); biases start at zero; Dense computes x.W + b with W
(in, out); inputs are cast to float32.
"""
import numpy as np
import torch
import torch.nn.functional as F


def layer_specs(board=19, input_dim=48, layers=12, filters_per_layer=128, filter_width_1=5, **kw):
    """[(filters, width)] for the k x k layers, following policy.py:81-122's kwargs."""
    out = []
    for i in range(1, layers + 1):
        width = kw.get("filter_width_%d" % i, filter_width_1 if i == 1 else 3)
        nb = kw.get("filters_per_layer_%d" % i, filters_per_layer)
        out.append((int(nb), int(width)))
    return out


def uniform_weights(specs, input_dim, board, head="policy", dense_units=256, seed=0):
    """Keras get_weights()-ordered list with init='uniform' kernels and zero biases."""
    rng = np.random.RandomState(seed)
    u = lambda *shape: rng.uniform(-0.05, 0.05, size=shape).astype(np.float32)  # noqa: E731
    w, cin = [], input_dim
    for nb, k in specs:
        w += [u(nb, cin, k, k), np.zeros(nb, np.float32)]
        cin = nb
    w += [u(1, cin, 1, 1), np.zeros(1, np.float32)]
    P = board * board
    if head == "policy":
        w += [np.zeros(P, np.float32)]
    else:
        w += [u(P, dense_units), np.zeros(dense_units, np.float32), u(dense_units, 1), np.zeros(1, np.float32)]
    return w


def forward(weights, x, specs, head="policy", kernel_flip=True, threads=None, return_logits=False):
    """x: (N, F, S, S) array -> (N, S*S) float32 probabilities, or (N,) values."""
    if threads:
        torch.set_num_threads(int(threads))
    with torch.no_grad():
        h = torch.as_tensor(np.asarray(x, dtype=np.float32))
        i = 0
        for nb, k in specs:
            W = torch.from_numpy(np.asarray(weights[i], np.float32))
            b = torch.from_numpy(np.asarray(weights[i + 1], np.float32))
            if kernel_flip:
                W = W.flip(2, 3)
            h = F.relu(F.conv2d(h, W, b, padding=k // 2))
            i += 2
        W = torch.from_numpy(np.asarray(weights[i], np.float32))
        b = torch.from_numpy(np.asarray(weights[i + 1], np.float32))
        h = F.conv2d(h, W, b)                      # 1x1: flipping is the identity
        i += 2
        flat = h.reshape(h.shape[0], -1)           # Flatten of (N,1,S,S): index x*S+y
        if head == "policy":
            logits = flat + torch.from_numpy(np.asarray(weights[i], np.float32))
            if return_logits:
                return logits.numpy()
            return torch.softmax(logits, dim=1).numpy()
        d1 = flat @ torch.from_numpy(np.asarray(weights[i], np.float32)) + torch.from_numpy(np.asarray(weights[i + 1], np.float32))
        d2 = d1 @ torch.from_numpy(np.asarray(weights[i + 2], np.float32)) + torch.from_numpy(np.asarray(weights[i + 3], np.float32))
        return torch.tanh(d2).reshape(-1).numpy()


def forward_resnet(weights, x, specs, unit_end, kernel_flip=True, threads=None, eps=1e-3):
    """ResnetPolicy.create_network (policy.py:141-271) at inference.  specs = [(filters, width)] of
    every convolution (the first one included), unit_end[i] = 1 where convolution i closes a
    residual unit.  weights in Keras get_weights() order: conv1 W, b; then per convolution
    BatchNormalization gamma, beta, running_mean, running_std -- each of shape (S,), because
    BatchNormalization() defaults to axis=-1 in Keras 1.2.0, the y coordinate under 'th'
    ordering -- and W, b; then the 1x1 W, b and the Bias vector."""
    if threads:
        torch.set_num_threads(int(threads))
    t = lambda a: torch.from_numpy(np.asarray(a, np.float32))  # noqa: E731
    with torch.no_grad():
        h = torch.as_tensor(np.asarray(x, dtype=np.float32))
        W, b = t(weights[0]), t(weights[1])
        k = specs[0][1]
        path = F.conv2d(h, W.flip(2, 3) if kernel_flip else W, b, padding=k // 2)          # linear
        block_input = path
        i = 2
        for c in range(1, len(specs)):
            gamma, beta, mean, var = (t(weights[i + j]) for j in range(4))
            path = (path - mean) * (gamma / torch.sqrt(var + eps)) + beta                   # broadcast over the last axis
            path = F.relu(path)
            W, b = t(weights[i + 4]), t(weights[i + 5])
            k = specs[c][1]
            path = F.conv2d(path, W.flip(2, 3) if kernel_flip else W, b, padding=k // 2)
            i += 6
            if unit_end[c]:
                path = block_input + path
                block_input = path
        path = F.relu(path)
        path = F.conv2d(path, t(weights[i]), t(weights[i + 1]))
        logits = path.reshape(path.shape[0], -1) + t(weights[i + 2])
        return torch.softmax(logits, dim=1).numpy()
