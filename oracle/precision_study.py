"""Which operand precision does the 13x192 policy need to keep the fp32 arg-max?  -- TEST INFRASTRUCTURE ONLY.

CPU emulation of the tensor-core arithmetic (16-bit operands, exact products, fp32 accumulation)
beside the fp32 forward of oracle/cnn_oracle.py (policy.py:94-138 / nn_util.py:48-58 of the
reference), so that precision schedules can be compared without a GPU.  Per layer a mode:

  h   fp16 activations x fp16 weights                       K x 1
  b   bf16 x bf16                                           K x 1
  a   (x_hi + x_lo) . W_hi        activations split         K x 2
  w   x_hi . (W_hi + W_lo)        weights split             K x 2
  x   x_hi.W_hi + x_lo.W_hi + x_hi.W_lo                     K x 3
  f   fp32 (no rounding)

plus `--diffuse`: weights rounded to fp16 with error feedback along the input channels so that the
rounding errors of every (filter, tap) sum to ~0 (ReLU activations have a large common mean).

    python -m oracle.precision_study --n 2048 --seeds 7 --schedules hhhhhhhhhhhh xhhhhhhhhhhh ...
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import cnn_oracle, go_oracle  # noqa: E402


def r16(t, kind="h"):
    return t.to(torch.float16 if kind == "h" else torch.bfloat16).to(torch.float32)


def diffuse_round(W):
    """fp16 rounding of W (out, in, kh, kw) with error feedback along `in`: the running sum of
    rounding errors of every (out, tap) stays within half an ulp."""
    W = W.clone()
    out = torch.empty_like(W)
    carry = torch.zeros_like(W[:, 0])
    for c in range(W.shape[1]):
        target = W[:, c] + carry
        q = r16(target)
        # never move a weight by more than one ulp of its own rounding
        q0 = r16(W[:, c])
        ulp = torch.maximum((q0.abs() * 2.0 ** -10), torch.full_like(q0, 2.0 ** -24))
        q = torch.where((q - q0).abs() <= ulp, q, q0)
        out[:, c] = q
        carry = W[:, c] + carry - q
    return out


# Expected mean of every canonical feature plane, as csrc/conv.cu's kPlaneDensity (rank="device" restates the
# device's channel ranking: response of every filter to these means, bias included, propagated through ReLU)
PLANE_DENSITY = [0.26, 0.26, 0.48, 1.00, 0.003, 0.003, 0.003, 0.003, 0.003, 0.003, 0.003, 0.50,
                 0.052, 0.071, 0.085, 0.074, 0.048, 0.040, 0.029, 0.117,
                 0.457, 0.009, 0.002, 0.001, 0.0005, 0.0003, 0.0002, 0.0006,
                 0.013, 0.002, 0.001, 0.0006, 0.0004, 0.0003, 0.0002, 0.0007,
                 0.018, 0.047, 0.107, 0.155, 0.041, 0.041, 0.020, 0.041,
                 0.014, 0.007, 0.466, 0.0, 0.471, 0.50]


def device_response(W, b, mean_in):
    """float32 response of every filter to the mean input, computed like rag_net_load_weights does (double sums)."""
    sw = W.double().sum(dim=(2, 3))                                   # (out, in): sum over taps
    return (b.double() + (sw * mean_in.double()[None, :]).sum(1)).float()


def rank_channels(W_prev, mean_in, how):
    """Order of a layer's OUTPUT channels, most energetic first.  `static`: by the filter's response
    to the mean input (sum over taps of W . mean_in) -- known at load time."""
    resp = (W_prev.sum(dim=(2, 3)) * mean_in[None, :]).sum(1)
    return torch.argsort(resp, descending=True)


def emulated_logits(weights, x, specs, schedule, diffuse=False, kernel_flip=True, G=32, rank="static", G1=16):
    h = torch.as_tensor(np.asarray(x, dtype=np.float32))
    i = 0
    order = None
    with torch.no_grad():
        mean_in = h.mean(dim=(0, 2, 3)) if rank != 'uniform' else torch.full((h.shape[1],), 0.25)
        if rank == "device":
            mean_in = torch.tensor(PLANE_DENSITY[:h.shape[1]], dtype=torch.float32)
        for l, (nb, k) in enumerate(specs):
            mode = schedule[l]
            W = torch.from_numpy(np.asarray(weights[i], np.float32))
            b = torch.from_numpy(np.asarray(weights[i + 1], np.float32))
            if kernel_flip:
                W = W.flip(2, 3)
            if mode == "p":
                # partial split: only the G most energetic input channels get the lo terms
                if l == 0:
                    top = torch.argsort(mean_in, descending=True)[:G1]
                elif rank == "true":
                    top = torch.argsort((h ** 2).mean(dim=(0, 2, 3)), descending=True)[:G]
                else:
                    top = order[:G]
                a_hi, w_hi = r16(h), r16(W)
                a_lo, w_lo = r16(h - a_hi), r16(W - w_hi)
                y = F.conv2d(a_hi, w_hi, b, padding=k // 2)
                y = y + F.conv2d(a_lo[:, top], w_hi[:, top], None, padding=k // 2)
                y = y + F.conv2d(a_hi[:, top], w_lo[:, top], None, padding=k // 2)
                if rank == "device":
                    resp = device_response(W, b, mean_in)
                    order = torch.sort(resp, descending=True, stable=True).indices
                    mean_in = torch.relu(resp)
                else:
                    order = rank_channels(W, mean_in, rank)
                    mean_in = h.mean(dim=(0, 2, 3)) if rank == "calib" else torch.relu((W.sum(dim=(2, 3)) * mean_in[None, :]).sum(1))
                h = F.relu(y)
                i += 2
                continue
            if rank == "device":
                resp = device_response(W, b, mean_in)
                order = torch.sort(resp, descending=True, stable=True).indices
                mean_in_next = torch.relu(resp)
            else:
                order = rank_channels(W, mean_in, rank)
                mean_in_next = torch.relu((W.sum(dim=(2, 3)) * mean_in[None, :]).sum(1))
            if mode == "f":
                a_op, w_op = h, W
            elif mode == "b":
                a_op, w_op = r16(h, "b"), r16(W, "b")
            else:
                a_hi = r16(h)
                w_hi = diffuse_round(W) if (diffuse and mode in "ha") else r16(W)
                a_op, w_op = a_hi, w_hi
                if mode in "ax":
                    a_op = a_hi + r16(h - a_hi)
                if mode in "wx":
                    w_op = w_hi + r16(W - w_hi)
            if mode == "x":
                # drop the lo.lo term like the device does
                a_lo = r16(h - a_hi)
                w_lo = r16(W - w_hi)
                y = F.conv2d(a_hi, w_hi + w_lo, b, padding=k // 2) + F.conv2d(a_lo, w_hi, None, padding=k // 2)
            else:
                y = F.conv2d(a_op, w_op, b, padding=k // 2)
            h = F.relu(y)
            mean_in = mean_in_next
            i += 2
        W = torch.from_numpy(np.asarray(weights[i], np.float32))
        b = torch.from_numpy(np.asarray(weights[i + 1], np.float32))
        h = F.conv2d(h, W, b)
        i += 2
        return (h.reshape(h.shape[0], -1) + torch.from_numpy(np.asarray(weights[i], np.float32))).numpy()


def stats(got, want):
    top_g, top_w = got.argmax(1), want.argmax(1)
    cg = got - got.mean(1, keepdims=True)
    cw = want - want.mean(1, keepdims=True)
    spread = want.max(1) - want.min(1)
    rel = np.sqrt(((cg - cw) ** 2).mean(1)) / spread
    srt = np.sort(want, 1)
    gap = (srt[:, -1] - srt[:, -2]) / spread
    n = len(got)
    p = float((top_g == top_w).mean())
    return {"top1": p, "miss": int((top_g != top_w).sum()), "n": n, "ci95": 1.96 * float(np.sqrt(max(p * (1 - p), 1e-9) / n)),
            "rms_rel": float(rel.mean()), "gap_median": float(np.median(gap))}


def cost(schedule, specs, cin0=48):
    """MMA work relative to a 3x3 192->192 layer (input channels padded to 64)."""
    unit = 9 * 192
    tot, cin = 0.0, cin0
    for (nb, k), m in zip(specs, schedule):
        kk = k * k * ((cin + 63) // 64 * 64)
        tot += kk / unit * {"h": 1, "b": 1, "a": 2, "w": 2, "x": 3, "f": 0, "p": 4 / 3}[m] * nb / 192
        cin = nb
    return tot


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=2048)
    ap.add_argument("--seeds", type=int, nargs="+", default=[7])
    ap.add_argument("--pos-seed", type=int, default=23)
    ap.add_argument("--layers", type=int, default=12)
    ap.add_argument("--filters", type=int, default=192)
    ap.add_argument("--schedules", nargs="+", default=["h" * 12])
    ap.add_argument("--diffuse", action="store_true")
    ap.add_argument("--G", type=int, default=32)
    ap.add_argument("--rank", default="static", choices=["static", "true", "calib", "uniform", "device"])
    ap.add_argument("--threads", type=int, default=os.cpu_count())
    a = ap.parse_args()
    torch.set_num_threads(a.threads)
    b, ag, ko, pl = go_oracle.random_playouts(a.n, size=19, seed=a.pos_seed)
    planes = go_oracle.features(b, ag, ko, pl, go_oracle.DEFAULT_FEATURES)
    specs = cnn_oracle.layer_specs(input_dim=48, layers=a.layers, filters_per_layer=a.filters)
    for seed in a.seeds:
        weights = cnn_oracle.uniform_weights(specs, 48, 19, seed=seed)
        want = cnn_oracle.forward(weights, planes, specs, return_logits=True)
        for sch in a.schedules:
            got = emulated_logits(weights, planes, specs, sch, diffuse=a.diffuse, G=a.G, rank=a.rank)
            st = stats(got, want)
            st.update(schedule=sch, seed=seed, cost=round(cost(sch, specs), 2), diffuse=a.diffuse, G=a.G, rank=a.rank)
            print(json.dumps(st), flush=True)


if __name__ == "__main__":
    main()
