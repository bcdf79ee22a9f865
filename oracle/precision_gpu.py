"""Top-1 agreement and speed of every operand precision of the device forward -- TEST INFRASTRUCTURE ONLY.

Runs the 13x192 policy (policy.py:94-138 of the reference) with Keras-'uniform' random weights on
random-playout positions in every dtype the library offers and compares the arg-max with the fp32
CPU forward of oracle/cnn_oracle.py (and, for all positions, with the device's own fp16x3 mode,
which the tests hold to the oracle).  Needs a GPU:

    python -m oracle.precision_gpu --n 8192 --seeds 7 8 9 --out gpurun_out/precision.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import cnn_oracle, go_oracle  # noqa: E402
from oracle.precision_study import stats  # noqa: E402


def device_logits(net, batch, chunk=4096):
    """fp32 output of the 1x1 convolution (read from the workspace of rag_eval_batch), chunk by chunk."""
    from rocalphago_b200 import BatchedGameState, _cabi
    L = _cabi.lib()
    outs = []
    P = batch.size * batch.size
    for i in range(0, len(batch), chunk):
        sub = BatchedGameState(batch.packed[i:i + chunk].contiguous(), batch.size)
        n = len(sub)
        net.model.eval_batch_device(sub, net.preprocessor.plane_src)
        off = (int(L.rag_features_workspace_bytes(n)) + 1023) // 1024 * 1024 + int(L.rag_net_logits_offset(net.model._handle, n))
        outs.append(net.model._ws[off:off + 4 * n * P].view(torch.float32).reshape(n, P).cpu().numpy())
    return np.concatenate(outs)


def layer_ms(net, batch, n=4096, reps=5):
    import ctypes
    from rocalphago_b200 import _cabi
    L = _cabi.lib()
    from rocalphago_b200 import BatchedGameState
    sub = BatchedGameState(batch.packed[:n].contiguous(), batch.size)
    h = net.model._ensure_handle(sub.device)
    L.rag_net_set_profiling(h, 1)
    best = None
    for _ in range(reps):
        net.model.eval_batch_device(sub, net.preprocessor.plane_src)
        torch.cuda.synchronize()
        ms = (ctypes.c_float * len(net.model.specs))()
        _cabi.check(L.rag_net_layer_times(h, ms, len(ms)))
        tot = float(sum(ms))
        best = tot if best is None else min(best, tot)
    L.rag_net_set_profiling(h, 0)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=8192)
    ap.add_argument("--n-oracle", type=int, default=None, help="positions also run through the fp32 CPU forward (default: all)")
    ap.add_argument("--seeds", type=int, nargs="+", default=[7, 8, 9])
    ap.add_argument("--dtypes", nargs="+", default=["fp16", "bf16", "fp16c32", "fp16c64", "fp16x3"])
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from rocalphago_b200 import BatchedGameState
    from rocalphago_b200.models import CNNPolicy
    feats = go_oracle.DEFAULT_FEATURES
    b, ag, ko, pl = go_oracle.random_playouts(a.n, size=19, seed=23, max_ply=381, stride=5)
    batch = BatchedGameState.from_arrays(b, ag, ko, pl)
    n_or = a.n if a.n_oracle is None else min(a.n, a.n_oracle)
    planes = go_oracle.features(b[:n_or], ag[:n_or], ko[:n_or], pl[:n_or], feats)
    specs = cnn_oracle.layer_specs(input_dim=48, layers=12, filters_per_layer=192)
    rows = []
    for seed in a.seeds:
        weights = cnn_oracle.uniform_weights(specs, 48, 19, seed=seed)
        t0 = time.time()
        want = cnn_oracle.forward(weights, planes, specs, threads=os.cpu_count(), return_logits=True)
        t_or = time.time() - t0
        got = {}
        for dt in a.dtypes:
            net = CNNPolicy(feats, layers=12, filters_per_layer=192, dtype=dt)
            net.model.set_weights(weights)
            got[dt] = device_logits(net, batch)
            ms = layer_ms(net, batch, n=min(4096, a.n))
            row = {"dtype": dt, "seed": seed, "conv_all_ms_at_%d" % min(4096, a.n): ms, "oracle_s": round(t_or, 1)}
            row["vs_oracle"] = stats(got[dt][:n_or], want)
            rows.append(row)
            del net
        if "fp16x3" in got:
            for row in rows:
                if row["seed"] == seed:
                    row["vs_fp16x3"] = stats(got[row["dtype"]], got["fp16x3"])
        for row in rows:
            if row["seed"] == seed:
                print(json.dumps(row), flush=True)
    # pooled over seeds
    pooled = {}
    for dt in a.dtypes:
        miss = sum(r["vs_oracle"]["miss"] for r in rows if r["dtype"] == dt)
        n = sum(r["vs_oracle"]["n"] for r in rows if r["dtype"] == dt)
        p = 1.0 - miss / n
        pooled[dt] = {"top1": p, "miss": miss, "n": n, "ci95": 1.96 * (max(p * (1 - p), 1e-12) / n) ** 0.5,
                      "conv_all_ms": float(np.mean([r[k] for r in rows if r["dtype"] == dt for k in r if k.startswith("conv_all_ms")]))}
    print(json.dumps({"pooled_vs_oracle": pooled}), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump({"rows": rows, "pooled_vs_oracle": pooled}, f, indent=1)


if __name__ == "__main__":
    main()
