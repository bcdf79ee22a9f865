"""The two CPU restatements of the reference's network forward agree with each other and with the
committed outputs on the reference's own network fixture.

oracle/cnn_oracle.py (torch conv2d with flipped kernels) and oracle/cnn_oracle2.py (scipy.signal.convolve2d,
the textbook true convolution Theano documents; numpy softmax / Dense / BatchNormalization) share no code.
Neither can be checked against Keras 1.2.0 itself here (not installable), which is what "parity unpinned"
in their headers means; this file pins everything short of that."""
import os

import numpy as np

from oracle import cnn_oracle, cnn_oracle2, go_oracle
from tests import golden_util as gu


def _minimodel():
    z = np.load(os.path.join(gu.GOLDEN, "minimodel.npz"))
    feats = [str(s) for s in z["feature_list"]]
    convs = [tuple(int(v) for v in c) for c in z["convs"]]
    return feats, convs, [z["w%02d" % i] for i in range(13)]


def test_minimodel_golden_outputs_are_reproduced_by_both_restatements():
    feats, convs, weights = _minimodel()
    gold = np.load(os.path.join(gu.GOLDEN, "minimodel_forward.npz"))
    g = gu.load_features("features_sgf.npz")
    planes = gu.select_planes(g["planes"], feats)[gold["index"]].astype(np.float32)
    assert np.array_equal(g["extra"]["ply"][gold["index"]], gold["ply"])
    p1 = cnn_oracle.forward(weights, planes, convs[:-1])
    assert np.abs(p1 - gold["probs"]).max() <= 1e-7
    sub = slice(0, 8)
    p2 = cnn_oracle2.forward(weights, planes[sub], convs[:-1])
    assert np.abs(p2 - gold["probs"][sub]).max() <= 1e-7
    assert np.allclose(gold["probs"].sum(1), 1.0, atol=1e-6)


def test_restatements_agree_on_the_13x192_policy():
    """BASELINE's network, Keras-'uniform' weights, one random-playout position (the scipy loop makes
    442k convolve2d calls per position)."""
    b, a, ko, pl = go_oracle.random_playouts(1, size=19, seed=77)
    planes = go_oracle.features(b, a, ko, pl, go_oracle.DEFAULT_FEATURES)
    specs = cnn_oracle.layer_specs(input_dim=48, layers=12, filters_per_layer=192)
    weights = cnn_oracle.uniform_weights(specs, 48, 19, seed=3)
    p1 = cnn_oracle.forward(weights, planes, specs)
    p2 = cnn_oracle2.forward(weights, planes, specs)
    assert np.abs(p1 - p2).max() <= 1e-6 * p2.max()
    assert p1.argmax() == p2.argmax()


def test_restatements_agree_on_value_head_and_other_shapes():
    rng = np.random.RandomState(5)
    x = rng.normal(0, 1, size=(3, 7, 9, 9)).astype(np.float32)
    specs = [(12, 5), (9, 1), (10, 7), (8, 3)]
    w = cnn_oracle.uniform_weights(specs, 7, 9, head="value", dense_units=11, seed=2)
    for i in range(1, len(w), 2):
        w[i] = rng.normal(0, 0.1, size=w[i].shape).astype(np.float32)          # non-zero biases
    v1 = cnn_oracle.forward(w, x, specs, head="value")
    v2 = cnn_oracle2.forward(w, x, specs, head="value")
    assert np.abs(v1 - v2).max() <= 1e-6
    wp = [(6.0 * a).astype(np.float32) for a in cnn_oracle.uniform_weights(specs, 7, 9, head="policy", seed=4)]
    wp[-1] = rng.normal(0, 0.3, size=81).astype(np.float32)                     # Bias layer
    p1, p2 = cnn_oracle.forward(wp, x, specs), cnn_oracle2.forward(wp, x, specs)
    assert np.abs(p1 - p2).max() <= 1e-7
    # the flip matters: cross-correlation (TensorFlow semantics) is a different function
    p3 = cnn_oracle.forward(wp, x, specs, kernel_flip=False)
    assert np.abs(p1 - p3).max() > 1e-4


def test_restatements_agree_on_the_residual_network_with_batchnorm_over_y():
    """ResnetPolicy (policy.py:141-271): BatchNormalization() with Keras-1 defaults normalises over the
    LAST axis -- the y coordinate under 'th' ordering -- so its four vectors have length S."""
    rng = np.random.RandomState(11)
    S, F, C = 9, 5, 8
    specs = [(C, 5), (C, 3), (C, 3), (C, 3), (C, 3)]
    unit_end = [1, 0, 1, 1, 1]
    x = rng.normal(0, 1, size=(2, F, S, S)).astype(np.float32)
    w = [rng.uniform(-0.2, 0.2, size=(C, F, 5, 5)).astype(np.float32), rng.normal(0, 0.1, size=C).astype(np.float32)]
    for _ in specs[1:]:
        w += [rng.uniform(0.5, 1.5, size=S).astype(np.float32), rng.normal(0, 0.2, size=S).astype(np.float32),
              rng.normal(0, 0.3, size=S).astype(np.float32), rng.uniform(0.5, 2.0, size=S).astype(np.float32)]
        w += [rng.uniform(-0.2, 0.2, size=(C, C, 3, 3)).astype(np.float32), rng.normal(0, 0.1, size=C).astype(np.float32)]
    w += [rng.uniform(-0.3, 0.3, size=(1, C, 1, 1)).astype(np.float32), np.zeros(1, np.float32), rng.normal(0, 0.3, size=S * S).astype(np.float32)]
    p1 = cnn_oracle.forward_resnet(w, x, specs, unit_end)
    p2 = cnn_oracle2.forward_resnet(w, x, specs, unit_end)
    assert np.abs(p1 - p2).max() <= 1e-6
    # a BatchNormalization over channels (axis 1) would be another function: S != C makes the shapes differ,
    # and transposing x <-> y must move the normalisation with it
    xt = np.ascontiguousarray(x.transpose(0, 1, 3, 2))
    p3 = cnn_oracle2.forward_resnet(w, xt, specs, unit_end)
    assert np.abs(p3.reshape(2, S, S).transpose(0, 2, 1).reshape(2, -1) - p2).max() > 1e-4
