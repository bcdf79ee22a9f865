"""Ad-hoc numerics check of the tcgen05 convolution stack against the CPU oracle (GPU box)."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from oracle import cnn_oracle, go_oracle
from rocalphago_b200.models import CNNPolicy, CNNValue
from rocalphago_b200 import BatchedGameState


def centered(logp):
    return logp - logp.mean(axis=1, keepdims=True)


def check(name, feats, n, board=19, dtype="bf16", head="policy", **arch):
    b, a, ko, pl = go_oracle.random_playouts(n, size=board, seed=11, max_ply=board * board, stride=5)
    planes = go_oracle.features(b, a, ko, pl, feats)
    cls = CNNPolicy if head == "policy" else CNNValue
    net = cls(feats, board=board, dtype=dtype, seed=3, **arch)
    w = net.model.get_weights()
    # sharpen: scale the 1x1 head so logits are O(1) and errors are visible through the softmax
    t0 = time.time()
    want = cnn_oracle.forward(w, planes, net.model.specs, head=head)
    tcpu = time.time() - t0
    torch.cuda.synchronize()
    try:
        got = net.forward(planes.astype(np.float32)).reshape(want.shape)
    except Exception as e:  # noqa: BLE001
        print("%-28s FAILED: %s" % (name, e))
        return
    torch.cuda.synchronize()
    if head == "policy":
        cg, cw = centered(np.log(got + 1e-30)), centered(np.log(want))
        rel = np.abs(cg - cw).max() / (cw.std() + 1e-12)
        top1 = (got.argmax(1) == want.argmax(1)).mean()
        print("%-28s n=%d max|dp|=%.3g  centered-logit err/std=%.4f  top1=%.4f  sum=%.6f  (cpu %.2fs)" % (
            name, n, np.abs(got - want).max(), rel, top1, got.sum(1).mean(), tcpu))
    else:
        print("%-28s n=%d max|dv|=%.3g  (value range %.3g..%.3g) (cpu %.2fs)" % (
            name, n, np.abs(got - want).max(), want.min(), want.max(), tcpu))
    # the fused device path must agree with forward(planes)
    batch = BatchedGameState.from_arrays(b, a, ko, pl)
    fused = net.model.eval_batch_device(batch, net.preprocessor.plane_src).cpu().numpy().reshape(got.shape)
    print("%-28s fused vs forward max diff %.3g" % ("", np.abs(fused - got).max()))


if __name__ == "__main__":
    F3 = ["board", "ones", "turns_since"]
    check("L1 f64 w3 (12 planes)", F3, 8, layers=1, filters_per_layer=64, filter_width_1=3)
    check("L1 f64 w5", F3, 8, layers=1, filters_per_layer=64)
    check("L2 f64", F3, 8, layers=2, filters_per_layer=64)
    check("L3 f32 48 planes", go_oracle.DEFAULT_FEATURES, 16, layers=3, filters_per_layer=32)
    check("L3 f32 fp16", go_oracle.DEFAULT_FEATURES, 16, dtype="fp16", layers=3, filters_per_layer=32)
    check("L12 f192 bf16", go_oracle.DEFAULT_FEATURES, 64, layers=12, filters_per_layer=192)
    check("L12 f192 fp16", go_oracle.DEFAULT_FEATURES, 64, dtype="fp16", layers=12, filters_per_layer=192)
    check("L12 f128 board13", go_oracle.DEFAULT_FEATURES, 32, board=13, layers=12, filters_per_layer=128)
    check("value L12 f152", go_oracle.DEFAULT_FEATURES + ["color"], 32, head="value")
    check("value L12 f192 fp16", go_oracle.DEFAULT_FEATURES + ["color"], 32, head="value", dtype="fp16", filters_per_layer=192)
