#!/usr/bin/env python
"""bench.py -- positions/sec of the batched position-evaluation hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--dtype fp16|bf16]
    python bench.py --impl reference ...        # the CPU path, for the driver's ratio

One step = one pass of the hot path over one batch of synthetic random-playout positions
(BASELINE.json: 48-plane features incl. ladders + 13-layer x 192-filter policy forward, 19x19):
packed states -> analyze/ladder/emit kernels -> 12 tcgen05 convolutions (+ fused 1x1) -> softmax.

  value     positions/s with the packed states already resident in HBM (device-timed, CUDA events)
  e2e       the same through the C-ABI with HOST buffers: pinned packed states -> H2D ->
            rag_eval_batch -> D2H of the probabilities, every step
  roofline  the dominant kernel (the 3x3 192->192 convolution, 11 of the 12 launches per step):
            algorithmic FLOPs per launch / its CUDA-event time, vs the measured dense bf16 peak
  cpu_baseline  the CPU port of the reference path (C oracle for the features on all host cores
            + torch fp32 forward), on a bounded sample

Multi-GPU (torchrun, one rank per GPU): every rank evaluates its own batch, no collective on the
data path ("scaling": "weak"); the timed region is bracketed by barriers and the slowest rank's
time is used.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "positions/sec: 48-plane features + policy CNN fwd, 19x19"
FLOP_3X3 = 2 * 361 * 192 * 192 * 9            # per position, one 3x3 192->192 layer
FLOP_POLICY = 2_801_452_416                    # per position, whole 13x192 policy
ISSUED = {"fp16": 1.0, "bf16": 1.0, "fp16c32": 4.0 / 3.0, "fp16c64": 5.0 / 3.0, "fp16x3": 3.0}   # MMA work per algorithmic flop, 3x3 layers
FEATURE_BYTES = 17_568                         # per position, packed state in + 48x361 u8 out


def ncu_traffic(key, n, algorithmic_bytes):
    """roofline.traffic: dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel, from the
    `ncu --set full` capture tools/ncu_extract.py summarised into profiles/ncu_traffic.json, scaled to n."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        e = json.load(open(path))[key]
    except (OSError, KeyError, ValueError):
        return None
    return {"bytes": int(e["bytes"] * n / e["positions_per_launch"]), "algorithmic_bytes": int(algorithmic_bytes),
            "source": e["source"] + " via profiles/ncu_traffic.json"}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], tf_burst=p["bf16_tflops"], tf_sustained=p["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback")


class ClockSampler(object):
    """nvidia-smi clocks and throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:  # noqa: BLE001
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------- CPU arm

def _cpu_features_worker(args):
    from oracle import go_oracle
    b, a, ko, pl = args
    return go_oracle.features(b, a, ko, pl, go_oracle.DEFAULT_FEATURES)


def cpu_path(n_positions, cores, seed=99, pool=None, weights=None, keep=False):
    """The reference's serial features -> forward (policy.py:42-45) in its CPU port: C oracle
    feature planes on `cores` processes, then the torch fp32 forward on `cores` threads.
    Returns (seconds_features, seconds_forward) and, with keep=True, the positions and the
    probabilities (the checker's side of the bench line's `parity` entry)."""
    import multiprocessing as mp
    from oracle import cnn_oracle, go_oracle
    b, a, ko, pl = go_oracle.random_playouts(n_positions, size=19, seed=seed)
    specs = cnn_oracle.layer_specs(input_dim=48, layers=12, filters_per_layer=192)
    if weights is None:
        weights = cnn_oracle.uniform_weights(specs, 48, 19, seed=1)
    chunks = [(b[i::cores], a[i::cores], ko[i::cores], pl[i::cores]) for i in range(cores) if len(b[i::cores])]
    own_pool = pool is None
    if own_pool:
        pool = mp.get_context("fork").Pool(cores)
        pool.map(_cpu_features_worker, [(c[0][:1], c[1][:1], c[2][:1], c[3][:1]) for c in chunks])  # spin up
    t0 = time.time()
    parts = pool.map(_cpu_features_worker, chunks)
    t1 = time.time()
    planes = np.concatenate(parts, axis=0)
    cnn_oracle.forward(weights, planes[:2], specs, threads=cores)
    t2 = time.time()
    probs = []
    for i in range(0, len(planes), 1024):               # bounded activations: 1,024 positions per forward call
        probs.append(cnn_oracle.forward(weights, planes[i:i + 1024], specs, threads=cores))
    t3 = time.time()
    if own_pool:
        pool.close()
    if keep:
        order = np.concatenate([np.arange(len(b))[i::cores] for i in range(cores)])     # the order the chunks were joined in
        return t1 - t0, t3 - t2, (b[order], a[order], ko[order], pl[order], np.concatenate(probs))
    return t1 - t0, t3 - t2



def _ref_python_worker(args):
    """Preprocess.state_to_tensor of the reference's own Python (oracle/_ref snapshot) on a few positions."""
    os.environ["RAG_REFERENCE_ROOT"] = os.path.join(ROOT, "oracle", "_ref")
    from oracle import ref_shim
    ref_shim.REFERENCE_ROOT = os.environ["RAG_REFERENCE_ROOT"]
    go, prep = ref_shim.load()
    b, a, ko, pl = args
    states = [ref_shim.state_from_arrays(go, b[i].astype(float), a[i], int(ko[i]), int(pl[i])) for i in range(len(b))]
    pp = prep.Preprocess(prep.DEFAULT_FEATURES)
    t0 = time.time()
    out = [pp.state_to_tensor(st) for st in states]
    return time.time() - t0, np.concatenate(out).astype(np.uint8)


def reference_python_features(n_single, n_per_core, cores, seed=4242, size=19, positions=None):
    """The reference's own feature code (AlphaGo/preprocessing/preprocessing.py:266-294 on AlphaGo/go.py),
    unmodified, timed on this host: one process on `n_single` random-playout positions, then one process per
    core on `n_per_core` positions each.  None where oracle/_ref is absent (python -m oracle.make_ref_snapshot
    writes it in the build container; it travels to the GPU box with the snapshot)."""
    from oracle import go_oracle, make_ref_snapshot
    if not make_ref_snapshot.available():
        return None
    import multiprocessing as mp
    if positions is None:
        # a random subset of many games' positions (consecutive ones would all come from one or two games)
        pool_b, pool_a, pool_ko, pool_pl = go_oracle.random_playouts(4096, size=size, seed=seed)
        pick = np.random.RandomState(seed).permutation(4096)[:n_single + n_per_core * cores]
        positions = pool_b[pick], pool_a[pick], pool_ko[pick], pool_pl[pick]
    b, a, ko, pl = positions
    t1, planes = _ref_python_worker((b[:n_single], a[:n_single], ko[:n_single], pl[:n_single]))
    want = go_oracle.features(b[:n_single], a[:n_single], ko[:n_single], pl[:n_single], go_oracle.DEFAULT_FEATURES)
    out = {"single": {"value": n_single / t1, "unit": "positions/s (features only)", "cores": 1,
                      "sample": "%d positions, %.2f s" % (n_single, t1),
                      "planes_equal_c_port": bool(np.array_equal(planes, want))}}
    if n_per_core > 0:
        chunks = [(b[n_single + i * n_per_core:n_single + (i + 1) * n_per_core], a[n_single + i * n_per_core:n_single + (i + 1) * n_per_core],
                   ko[n_single + i * n_per_core:n_single + (i + 1) * n_per_core], pl[n_single + i * n_per_core:n_single + (i + 1) * n_per_core])
                  for i in range(cores)]
        with mp.get_context("fork").Pool(cores) as pool:
            pool.map(_ref_python_worker, [(c[0][:1], c[1][:1], c[2][:1], c[3][:1]) for c in chunks])     # imports, spin-up
            t0 = time.time()
            pool.map(_ref_python_worker, chunks)
            tall = time.time() - t0
        out["all_cores"] = {"value": n_per_core * cores / tall, "unit": "positions/s (features only)", "cores": cores,
                            "sample": "%d positions on %d processes, %.2f s" % (n_per_core * cores, cores, tall)}
    return out


def reference_entries(cores, forward_pos_per_s):
    """cpu_baseline entries for the reference's own Python: its features (1 core / all cores) followed serially
    by the fp32 forward (policy.py:42-45 runs them one after the other; Keras cannot run here, so the forward
    is the torch restatement at `forward_pos_per_s`)."""
    try:
        ref = reference_python_features(24, 4, cores)
    except Exception as e:  # noqa: BLE001 -- a reported baseline must never cost the bench line
        return {"reference_error": str(e)[:200]}
    if ref is None:
        return {}
    out = {}
    for key, name in (("single", "reference"), ("all_cores", "reference_all_cores")):
        f = ref[key]["value"]
        out[name] = {"value": 1.0 / (1.0 / f + 1.0 / forward_pos_per_s), "unit": "positions/s", "cores": ref[key]["cores"],
                     "kind": "reference", "features_only": f,
                     "sample": "Preprocess.state_to_tensor of the unmodified reference (oracle/_ref): " + ref[key]["sample"]
                               + "; then the torch fp32 forward at %.0f pos/s" % forward_pos_per_s}
    out["reference"]["planes_equal_c_port"] = ref["single"]["planes_equal_c_port"]
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sample = args.ref_sample
    import multiprocessing as mp
    pool = mp.get_context("fork").Pool(cores)
    for _ in range(args.warmup):
        cpu_path(min(sample, 64), cores, pool=pool)
    t0 = time.time()
    tf = tn = 0.0
    for k in range(args.steps):
        a, b = cpu_path(sample, cores, seed=1000 + k, pool=pool)
        tf += a; tn += b
    total = time.time() - t0
    pool.close()
    value = sample * args.steps / (tf + tn)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "positions/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * (tf + tn) / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "13x192 policy + 48 planes, 19x19; CPU port of the reference path "
                               "(C feature oracle on all cores, then torch fp32 forward), %d positions/step" % sample,
                   "batch_per_gpu": sample},
        "cpu_baseline": {"value": value, "unit": "positions/s", "cores": cores, "kind": "port",
                         "sample": "%d steps x %d random-playout positions; features %.2fs + forward %.2fs; wall %.1fs"
                                   % (args.steps, sample, tf, tn, total)},
        "e2e": {"value": value, "unit": "positions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    line["cpu_baseline"].update(reference_entries(cores, sample * args.steps / tn))
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------- GPU arm

def run_gpu(args):
    import torch
    import torch.distributed as dist
    from rocalphago_b200 import BatchedGameState, _cabi
    from rocalphago_b200.models import CNNPolicy
    from rocalphago_b200.pipeline import EvalPipeline
    from rocalphago_b200.preprocessing import DEFAULT_FEATURES
    from rocalphago_b200.state import _ptr, _stream

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    n, K, W = args.batch, args.steps, max(3, args.warmup)
    trace = (lambda *a: print("[rank %d]" % rank, *a, file=sys.stderr, flush=True)) if os.environ.get("BENCH_TRACE") else (lambda *a: None)
    trace("process group up")
    L = _cabi.lib()
    net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype=args.dtype, seed=1234)
    if args.workload == "value":
        # BASELINE config 3: value net, 49 planes (48 + colour), 12 conv + 1x1 + Dense 256 + tanh
        from rocalphago_b200.models import CNNValue
        net = CNNValue(DEFAULT_FEATURES + ["color"], layers=12, filters_per_layer=192, dtype=args.dtype, seed=4321)
    vnet = None
    if args.workload == "policy+value":
        from rocalphago_b200.models import CNNValue
        vnet = CNNValue(DEFAULT_FEATURES + ["color"], layers=12, filters_per_layer=192, dtype=args.dtype, seed=4321)
    model = net.model
    plane_src = net.preprocessor.plane_src
    h = model._ensure_handle(dev)
    _cabi.check(L.rag_net_set_profiling(h, 1))
    L.rag_set_conv_pair(args.conv_pair)

    # synthetic inputs: `pool` distinct batches of random-playout positions generated on the device
    pool = max(2, args.pool)
    batches = [BatchedGameState.random_playouts(n, seed=20261106 + 7919 * (rank * pool + i), device=dev) for i in range(pool)]
    host_in = [b.packed.cpu().pin_memory() for b in batches]
    torch.cuda.synchronize()
    trace("inputs ready")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def layer_times():
        lm = (ctypes.c_float * 12)()
        _cabi.check(L.rag_net_layer_times(h, lm, 12))
        return np.asarray(list(lm))

    sampler = ClockSampler(local)
    serial_ms = None
    if args.serial:
        # one stream, rag_eval_batch per step, 256 MiB written between steps to flush L2
        out = torch.empty((n, 361), dtype=torch.float32, device=dev)
        ws_bytes = int(L.rag_eval_workspace_bytes(h, n))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        src_ptr = plane_src.ctypes.data_as(ctypes.c_void_p)

        def step(packed):
            _cabi.check(L.rag_eval_batch(h, _ptr(packed), n, src_ptr, len(plane_src), _ptr(out), _ptr(ws), ws_bytes, _stream()))
        for i in range(W):
            step(batches[i % pool].packed)
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for k in range(K):
            flush.zero_()
            ev[k][0].record()
            step(batches[k % pool].packed)
            ev[k][1].record()
        barrier()
        serial_ms = sum(a.elapsed_time(b) for a, b in ev) / K
        del ws, flush

    if args.no_pipeline:
        print(json.dumps({"serial_ms_per_step": serial_ms, "conv_pair": args.conv_pair}))
        return
    pipe = EvalPipeline(net, n, device=dev, value_net=vnet, slots=args.slots)
    first = (lambda r: r[0]) if vnet is not None else (lambda r: r)
    one = (lambda t: float(t.reshape(-1)[0]))
    for i in range(W):
        pipe.submit(batches[i % pool].packed)
    pipe.drain()
    trace("warm-up done")
    barrier()
    trace("barrier passed")
    if rank == 0:
        sampler.start()

    # ---- device-resident inputs: K batches through the two-stage pipeline
    layer_ms, samples = np.zeros(12), 0
    wall0 = time.time()
    e0 = pipe.mark()
    for k in range(K):
        pipe.submit(batches[k % pool].packed)
        if os.environ.get("BENCH_TRACE"):
            print("submitted", k, file=sys.stderr, flush=True)
        if K >= 8 and k in (K // 3, 2 * K // 3) and not os.environ.get("BENCH_NO_SAMPLE"):   # two samples of the per-layer events (each costs one pipeline bubble)
            layer_ms += layer_times(); samples += 1
    e1 = pipe.mark()
    pipe.drain()
    barrier()
    wall = time.time() - wall0
    dev_ms = e0.elapsed_time(e1)
    trace("device-resident pass done", dev_ms)
    if samples == 0:
        layer_ms, samples = layer_times(), 1
    layer_ms /= samples

    # ---- end to end with host buffers: pinned packed states -> H2D -> features -> net -> D2H, every step;
    #      the caller collects each result (on the host) as late as the pipeline allows (slots - 1 steps behind)
    #      -- through a pipeline that replays its stages as CUDA graphs (host cost per step 0.45 -> 0.06 ms, which
    #      matters when eight ranks share the host's cores); per-layer profiling events cannot be captured, so off
    _cabi.check(L.rag_net_set_profiling(h, 0))
    if args.e2e_graphs:
        pipe.drain()
        pipe = EvalPipeline(net, n, device=dev, value_net=vnet, slots=args.slots, graphs=True)
    for i in range(2 * pipe.nslots + 1):
        pipe.submit(host_in[i % pool], to_host=True)
    pipe.drain()
    barrier()
    t0 = time.time()
    f0 = pipe.mark()
    checksum = 0.0
    lag = pipe.nslots - 1            # a result stays readable until `slots` later batches have been submitted
    for k in range(K):
        t = pipe.submit(host_in[k % pool], to_host=True)
        if k >= lag:
            checksum += one(first(pipe.result(t - lag)))
    for t in range(max(pipe.count - lag, pipe.count - K), pipe.count):
        checksum += one(first(pipe.result(t)))
    f1 = pipe.mark()
    pipe.drain()
    barrier()
    e2e_wall_ms = (time.time() - t0) * 1e3
    e2e_ms = f0.elapsed_time(f1)
    trace("e2e pass done", e2e_ms)
    clocks = sampler.stop() if rank == 0 else None

    # results gathered to every rank (NCCL all_gather over NVLink) -- off the data path, timed on its own
    gather_ms = None
    if world > 1:
        from rocalphago_b200 import distributed as D
        local = pipe.slots[(pipe.count - 1) % pipe.nslots].out
        D.gather_results(local, n * world)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        full = D.gather_results(local, n * world)
        g1.record()
        torch.cuda.synchronize()
        assert full.shape[0] == n * world
        gather_ms = g0.elapsed_time(g1)

    t = torch.tensor([dev_ms, e2e_ms, e2e_wall_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, e2e_wall_ms = [float(v) for v in t.cpu()]

    if rank == 0:
        peaks = load_peaks()
        total_pos = n * K * world
        value = total_pos / (dev_ms / 1e3)
        e2e_val = total_pos / (max(e2e_ms, e2e_wall_ms) / 1e3)
        k3 = float(np.mean(layer_ms[1:]))                      # the eleven 3x3 192->192 launches
        achieved = FLOP_3X3 * n / (k3 / 1e3) / 1e12
        act_mb = n * 361 * 192 * 2 * 2 / 1e6              # the two ping-pong activation buffers every layer reads and writes
        l2_note = ("no flush needed: every layer of a step reads and writes %.0f MB of activations (%.1fx the 126 MB L2) and the "
                   "input batches rotate; the --serial pass writes 256 MiB between steps" % (act_mb, act_mb / 126)) if act_mb > 252 else \
                  ("NOT flushed: a step's activations (%.0f MB) fit the 126 MB L2 at this batch size, as they do for a "
                   "service running such batches back to back; the headline configuration is batch 4096" % act_mb)
        line = {
            "metric": ("positions/sec: 49-plane features + value CNN fwd, 19x19" if args.workload == "value" else METRIC)
                      if vnet is None else "positions/sec: 48/49-plane features + policy AND value CNN fwd, 19x19",
            "value": value, "unit": "positions/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": dev_ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": ("%s, 19x19, batch %d/GPU, device-generated random-playout positions, random-init weights"
                                    % ({"policy": "48 planes (incl. ladders) + 13x192 policy forward",
                                        "value": "49 planes (48 incl. ladders + colour) + 13x192 value net forward (Dense 256, tanh)",
                                        "policy+value": "48/49 planes (one shared analysis) + 13x192 policy AND value forwards"}[args.workload], n)),
                       "batch_per_gpu": n,
                       "schedule": "two-stage pipeline, %d batches in flight: the feature kernels of the next batches run "
                                   "on their own streams underneath the convolutions of the current one "
                                   "(rocalphago_b200/pipeline.py)" % pipe.nslots,
                       "l2": l2_note,
                       "parallelism": "dp%d (independent batches, no collective on the data path)" % world},
            "e2e": {"value": e2e_val, "unit": "positions/s", "h2d_bytes_per_step": int(n * 768),
                    "d2h_bytes_per_step": (int(n * 4) if args.workload == "value" else int(n * 361 * 4) + (int(n * 4) if vnet is not None else 0)),
                    "ms_per_step": max(e2e_ms, e2e_wall_ms) / K,
                    "checksum": checksum},
            "gpu_launches": (16 if vnet is None else 30) * K,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "conv_umma_kernel, 3x3 192->192 layer (11 of 12 conv launches per step)",
                         "achieved": achieved, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["tf_sustained"], "peak_source": peaks["src"] + " (sustained)",
                         "ms_per_launch": k3,
                         # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture
                         # of this kernel at batch 4096 (profiles/r1_final_ncu_summary.md); scales with the batch
                         "traffic": ncu_traffic("conv_umma/" + args.dtype, n, n * 361 * 2 * 2 * {"fp16c32": 256, "fp16c64": 256, "fp16x3": 384}.get(args.dtype, 192)),
                         "issued_over_algorithmic": ISSUED[args.dtype],
                         "issued_tflops": achieved * ISSUED[args.dtype],
                         "frac_issued": achieved * ISSUED[args.dtype] / peaks["tf_sustained"],
                         "note": "CUDA events recorded by the library on the launching stream, sampled %d time(s) inside "
                                 "the timed region while the next batch's feature kernels share the SMs" % samples
                                 + ("; `achieved` counts ALGORITHMIC flops (2*361*192*192*9 per position): this precision issues "
                                    "%.2fx as many MMAs (extra K steps for the corrected channels), so the tensor pipe runs at "
                                    "`issued_tflops`" % ISSUED[args.dtype] if ISSUED[args.dtype] > 1 else "")},
            "stage_ms": {"conv_layer1": float(layer_ms[0]), "conv_3x3_mean": k3, "conv_all": float(layer_ms.sum()),
                         "pipelined_step": dev_ms / K, "serial_step": serial_ms},
            "whole_net_tflops": FLOP_POLICY * n / (float(layer_ms.sum()) / 1e3) / 1e12,
            "wall_s": wall,
            "gather_ms": gather_ms,
        }
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            sample = args.cpu_sample
            if args.workload == "policy":
                # the same positions and weights through the device path, checked against the CPU forward just timed
                tf, tn, (pb, pa, pko, ppl, want) = cpu_path(sample, cores, weights=model.get_weights(), keep=True)
                chk = BatchedGameState.from_arrays(pb, pa, pko, ppl, device=dev)
                got = np.concatenate([model.eval_batch_device(BatchedGameState(chk.packed[i:i + n].contiguous(), 19), plane_src).cpu().numpy()
                                      for i in range(0, sample, n)])
                miss = int((got.argmax(1) != want.argmax(1)).sum())
                line["parity"] = {"dtype": args.dtype, "top1": 1.0 - miss / float(sample), "miss": miss, "n": sample,
                                  "max_abs": float(np.abs(got - want).max()), "gate": {"top1": 0.995, "max_abs": 2e-3},
                                  "against": "the fp32 CPU forward timed for cpu_baseline, same positions and weights; "
                                             "outside the timed region (the 3-seed, 24,576-position gate is tests/test_gpu_network.py)"}
                # the reference-facing Python call on host objects: CNNPolicy.batch_eval_state(list of n GameState-likes)
                from rocalphago_b200 import PositionView
                views = [PositionView(pb[i].astype(np.float64), pa[i].astype(np.int64), int(pko[i]), int(ppl[i])) for i in range(min(n, sample))]
                best = None
                for _ in range(4):
                    t_py = time.perf_counter()
                    res = net.batch_eval_state(views)
                    first_state = res[0]
                    t_py = time.perf_counter() - t_py
                    best = t_py if best is None else min(best, t_py)
                line["e2e_python"] = {"value": len(views) / best, "unit": "positions/s", "ms_per_call": 1e3 * best, "states": len(views),
                                      "note": "CNNPolicy.batch_eval_state(list of GameState-like objects with float64 boards) -> "
                                              "MoveDistributions (the reference's list of lists of ((x, y), p), tuples built when a "
                                              "state is looked at); one synchronous call, best of 4; first state has %d moves" % len(first_state)}
            else:
                tf, tn = cpu_path(sample, cores)
            line["cpu_baseline"] = {"value": sample / (tf + tn), "unit": "positions/s", "cores": cores, "kind": "port",
                                    "sample": "%d random-playout positions once: C feature oracle on %d processes %.2fs "
                                              "(%.0f pos/s) + torch fp32 13x192 forward on %d threads %.2fs (%.0f pos/s)"
                                              % (sample, cores, tf, sample / tf, cores, tn, sample / tn)}
            line["cpu_baseline"].update(reference_entries(cores, sample / tn))
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_features(args):
    """BASELINE config 4: feature extraction only (all 48 planes incl. ladders) -> uint8 NCHW, over
    `steps` x `batch` DISTINCT device-generated random-playout positions per GPU (default 32 x 32,768 =
    1,048,576; never recycled inside the timed region when steps <= 32).
    Roofline: the emit kernel, the HBM-bound kernel of the feature path (17,568 B per position)."""
    import torch
    import torch.distributed as dist
    from rocalphago_b200 import BatchedGameState, Preprocess, _cabi
    from rocalphago_b200 import preprocessing as P
    from rocalphago_b200.state import _ptr
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n, K, W = args.batch, args.steps, max(3, args.warmup)
    L = _cabi.lib()
    pp = Preprocess()
    npool = min(K, 32)
    pool = [BatchedGameState.random_playouts(n, seed=20261106 + 7919 * (rank * 64 + i), device=dev) for i in range(npool)]
    # `lanes` batches in flight, each on its own stream with its own workspace and output buffer: the
    # tail of one batch's ladder search (a few warps) runs under the bulk of the next batch's
    lanes = max(1, args.slots or 4)
    streams = [torch.cuda.Stream(device=dev) for _ in range(lanes)]
    wss = [torch.empty(int(L.rag_features_workspace_bytes(n)), dtype=torch.uint8, device=dev) for _ in range(lanes)]
    outs = [torch.empty((n, 48, 19, 19), dtype=torch.uint8, device=dev) for _ in range(lanes)]
    src = np.ascontiguousarray(pp.plane_src, np.int32)
    vp = ctypes.c_void_p

    def step(packed, lane, ev=None):
        st = vp(streams[lane].cuda_stream)
        _cabi.check(L.rag_features_analyze(_ptr(packed), n, 19, 1, _ptr(wss[lane]), wss[lane].numel(), st))
        if ev: ev[0].record(streams[lane])
        _cabi.check(L.rag_features_emit(_ptr(wss[lane]), n, 19, src.ctypes.data_as(vp), len(src), P.LAYOUT_NCHW_U8, 0, 0,
                                        _ptr(outs[lane]), st))
        if ev: ev[1].record(streams[lane])
    for i in range(max(W, lanes)):
        step(pool[i % npool].packed, i % lanes)
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0: sampler.start()
    E = lambda: torch.cuda.Event(enable_timing=True)
    evs = [(E(), E()) for _ in range(K)]
    t0, t1 = E(), E()
    t0.record()
    for s_ in streams: s_.wait_event(t0)
    # the next batch goes to the lane with the fewest unfinished steps (at most two queued per lane): one batch with
    # an exceptionally long ladder search then delays the batches behind it on ITS lane only, not a fixed quarter of the sweep
    pending = [[] for _ in range(lanes)]
    for k in range(K):
        while True:
            for q in pending:
                while q and q[0].query():
                    q.pop(0)
            lane = min(range(lanes), key=lambda i: len(pending[i]))
            if len(pending[lane]) < 2:
                break
        step(pool[k % npool].packed, lane, evs[k])
        pending[lane].append(evs[k][1])
    for s_ in streams: torch.cuda.current_stream().wait_stream(s_)
    t1.record()
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    tot = t0.elapsed_time(t1)

    # ---- end to end with host buffers: pinned packed states -> H2D -> features -> D2H of the uint8 planes,
    #      every step, two lanes; the planes are 17,328 B per position, so this leg is bound by the host link
    e2e_steps = min(K, 8)
    h_in = [pool[i % npool].packed.cpu().pin_memory() for i in range(2)]
    h_out = [torch.empty((n, 48, 19, 19), dtype=torch.uint8).pin_memory() for _ in range(2)]
    d_in = [torch.empty_like(pool[0].packed) for _ in range(2)]

    def e2e_step(k):
        lane = k % 2
        with torch.cuda.stream(streams[lane]):
            d_in[lane].copy_(h_in[lane], non_blocking=True)
            step(d_in[lane], lane)
            h_out[lane].copy_(outs[lane], non_blocking=True)
    for k in range(2):
        e2e_step(k)
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    w0 = time.time()
    for k in range(e2e_steps):
        e2e_step(k)
    torch.cuda.synchronize()
    e2e_ms = (time.time() - w0) * 1e3
    checksum = int(h_out[0][0].sum()) + int(h_out[1][0].sum())
    if world > 1: dist.barrier()
    clocks = sampler.stop() if rank == 0 else None

    # the emit kernel alone (nothing else on the GPU), 256 MiB written before each launch: its roofline figure
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    solo = [(E(), E()) for _ in range(5)]
    with torch.cuda.stream(streams[0]):
        for a, b in solo:
            flush.zero_()
            a.record(streams[0])
            _cabi.check(L.rag_features_emit(_ptr(wss[0]), n, 19, src.ctypes.data_as(vp), len(src), P.LAYOUT_NCHW_U8, 0, 0,
                                            _ptr(outs[0]), vp(streams[0].cuda_stream)))
            b.record(streams[0])
    torch.cuda.synchronize()
    emit = sum(a.elapsed_time(b) for a, b in solo) / len(solo)
    # what a write-only kernel can reach on this part: a plain fill of the same buffer, timed the same way
    fills = [(E(), E()) for _ in range(5)]
    with torch.cuda.stream(streams[0]):
        for a, b in fills:
            flush.zero_()
            a.record(streams[0])
            outs[1 % lanes].zero_()
            b.record(streams[0])
    torch.cuda.synchronize()
    fill = sum(a.elapsed_time(b) for a, b in fills) / len(fills)
    t = torch.tensor([tot, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1: dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot, e2e_ms = [float(v) for v in t.cpu()]
    if rank == 0:
        peaks = load_peaks()
        achieved = FEATURE_BYTES * n / (emit / 1e3) / 1e9
        line = {
            "metric": "positions/sec: 48-plane feature extraction only, 19x19", "value": n * K * world / (tot / 1e3),
            "unit": "positions/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": tot / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "feature-extraction-only sweep (BASELINE config 4): 48 planes incl. ladders -> uint8 NCHW, "
                                   "%d steps x %d = %d positions per GPU, %d distinct device-generated random-playout batches"
                                   % (K, n, K * n, npool),
                       "batch_per_gpu": n, "positions_per_gpu": K * n,
                       "schedule": "%d lanes (streams), at most two steps queued per lane, the next batch goes to the least loaded lane" % lanes,
                       "l2": "no flush needed: every step writes %.0f MB of planes (%.1fx the 126 MB L2) and every batch is new"
                             % (n * 17328 / 1e6, n * 17328 / 126e6),
                       "parallelism": "dp%d (independent positions, no collective)" % world},
            "e2e": {"value": n * e2e_steps * world / (e2e_ms / 1e3), "unit": "positions/s", "h2d_bytes_per_step": int(n * 768),
                    "d2h_bytes_per_step": int(n * 17328), "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps, "checksum": checksum,
                    "note": "bound by the device-to-host copy of the planes (17,328 B per position over the host link)"},
            "gpu_launches": 3 * K, "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "emit_u8_flat_kernel", "achieved": achieved, "peak": peaks["hbm"], "unit": "GB/s",
                         "frac": achieved / peaks["hbm"], "ms_per_launch": emit, "traffic": ncu_traffic("emit_u8_flat/u8", n, FEATURE_BYTES * n),
                         "write_only": {"fill_gbs": n * 17328 / (fill / 1e3) / 1e9, "fill_ms": fill, "emit_over_fill": fill / emit,
                                        "note": "the kernel only writes (planes in: 0.2 of the bytes out); a plain fill of the same "
                                                "buffer is what this part's HBM takes for write-only traffic -- `peak` is a COPY rate "
                                                "(read + write), which a write-only kernel cannot reach"},
                         "peak_source": peaks["src"], "note": "17,568 algorithmic bytes per position (SURVEY 8d); the kernel timed alone after "
                         "the timed region (inside it, it overlaps other batches' ladder searches); the step itself is bound by "
                         "the latency of the ladder searches, not by HBM"},
            "stage_ms": {"step": tot / K, "emit_u8_alone": emit,
                         "analysis_and_ladders_mean": float(np.mean([a.elapsed_time(b) for a, b in evs]))}}
        if world == 1 and not args.no_cpu_baseline:
            import multiprocessing as mp
            from oracle import go_oracle
            cores = os.cpu_count() or 1
            sample = args.cpu_sample * 8
            b, a, ko, pl = go_oracle.random_playouts(sample, size=19, seed=99)
            chunks = [(b[i::cores], a[i::cores], ko[i::cores], pl[i::cores]) for i in range(cores)]
            with mp.get_context("fork").Pool(cores) as cp:
                cp.map(_cpu_features_worker, [(c[0][:1], c[1][:1], c[2][:1], c[3][:1]) for c in chunks])
                c0 = time.time()
                cp.map(_cpu_features_worker, chunks)
                tf = time.time() - c0
            line["cpu_baseline"] = {"value": sample / tf, "unit": "positions/s", "cores": cores, "kind": "port",
                                    "sample": "%d random-playout positions once: C feature oracle on %d processes, %.2f s" % (sample, cores, tf)}
            try:
                ref = reference_python_features(24, 4, cores)
            except Exception as e:  # noqa: BLE001
                ref, line["cpu_baseline"]["reference_error"] = None, str(e)[:200]
            if ref:
                line["cpu_baseline"]["reference"] = dict(ref["single"], kind="reference")
                line["cpu_baseline"]["reference_all_cores"] = dict(ref["all_cores"], kind="reference")
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_cfg1(args):
    """BASELINE config 1 (the reference's own CPU-runnable case): Preprocess(48 default features) +
    CNNPolicy(layers=3, filters_per_layer=32).batch_eval_state on 16 positions of the reference's SGFs (plies
    20 / 41 / 101 / 161 of the Lee Sedol games; states recorded from the reference's replay in
    tests/golden/features_sgf.npz), through the reference-facing Python call on GameState-like objects
    (policy.py:27-52) -- lists of ((x, y), p) out.  One step = one such call."""
    import torch
    from rocalphago_b200 import PositionView
    from rocalphago_b200.models import CNNPolicy
    from rocalphago_b200.preprocessing import DEFAULT_FEATURES
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(0)
    z = np.load(os.path.join(ROOT, "tests", "golden", "features_sgf.npz"))
    names = [str(v) for v in z["names"]]
    pick = [i for i in range(len(z["ply"])) if "Lee-Sedol" in names[int(z["game"][i])] and int(z["ply"][i]) in (20, 41, 101, 161)][:16]
    b, a, ko, pl = z["board"][pick], z["age"][pick], z["ko"][pick], z["player"][pick]
    states = [PositionView(b[i], a[i], None if ko[i] < 0 else (int(ko[i]) // 19, int(ko[i]) % 19), int(pl[i])) for i in range(len(pick))]
    net = CNNPolicy(DEFAULT_FEATURES, layers=3, filters_per_layer=32, seed=16)
    K, W = args.steps, max(3, args.warmup)
    for _ in range(W):
        res = net.batch_eval_state(states)
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    lat = []
    for _ in range(K):
        t0 = time.perf_counter()
        res = net.batch_eval_state(states)
        lat.append(time.perf_counter() - t0)
    clocks = sampler.stop()
    lat = np.asarray(lat)
    n = len(states)
    line = {"metric": METRIC, "value": n / float(lat.mean()), "unit": "positions/s", "n_gpus": 1, "steps": K, "warmup": W,
            "ms_per_step": 1e3 * float(lat.mean()), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": net.model.dtype, "data": "16 positions of the reference's SGF fixtures, random-init weights",
            "config": {"workload": "BASELINE config 1: 48 planes + CNNPolicy(layers=3, filters_per_layer=32).batch_eval_state, "
                                   "16 SGF positions as GameState-like objects, lists of ((x, y), p) out", "batch_per_gpu": n,
                       "l2": "not flushed: a 16-position step touches < 1 MB; this configuration measures call latency"},
            "e2e": {"value": n / float(lat.mean()), "unit": "positions/s", "h2d_bytes_per_step": n * 768, "d2h_bytes_per_step": n * 361 * 4,
                    "ms_per_step": 1e3 * float(lat.mean()), "ms_p50": 1e3 * float(np.median(lat)), "ms_p99": 1e3 * float(np.percentile(lat, 99)),
                    "note": "host objects in, Python lists out, every step: value and e2e are the same measurement here"},
            "gpu_launches": 7 * K, "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "conv_umma_kernel (3 launches per step)", "achieved": 41_055_808 * n / float(lat.mean()) / 1e12,
                         "peak": load_peaks()["tf_sustained"], "unit": "TFLOP/s", "frac": 41_055_808 * n / float(lat.mean()) / 1e12 / load_peaks()["tf_sustained"],
                         "traffic": None, "note": "41.06 MFLOP per position (SURVEY 8d); a 16-position call is bound by launch and "
                         "host latency, not by the tensor pipe -- the fraction is reported for completeness"}}
    if not args.no_cpu_baseline:
        from oracle import cnn_oracle, go_oracle
        cores = os.cpu_count() or 1
        planes = go_oracle.features(b, a, ko, pl, go_oracle.DEFAULT_FEATURES)
        want = cnn_oracle.forward(net.model.get_weights(), planes, net.model.specs, threads=cores)
        got = np.zeros_like(want)
        for i, r in enumerate(res):
            for (x, y), p in r:
                got[i, x * 19 + y] = p
        legal = got > 0
        ref_norm = np.where(legal, want, 0)
        ref_norm = ref_norm / ref_norm.sum(1, keepdims=True)
        line["parity"] = {"max_abs": float(np.abs(got - ref_norm).max()), "top1": float((got.argmax(1) == ref_norm.argmax(1)).mean()), "n": n,
                          "against": "C feature oracle + fp32 CPU forward, renormalised over the returned moves"}
        t0 = time.time()
        for _ in range(20):
            cnn_oracle.forward(net.model.get_weights(), planes, net.model.specs, threads=cores)
        t_fwd = (time.time() - t0) / 20
        ref = reference_python_features(n, 0, cores, positions=(b, a, ko, pl))
        t0 = time.time()
        for _ in range(20):
            go_oracle.features(b, a, ko, pl, go_oracle.DEFAULT_FEATURES)
        t_port = (time.time() - t0) / 20
        line["cpu_baseline"] = {"value": n / (t_port + t_fwd), "unit": "positions/s", "cores": cores, "kind": "port",
                                "sample": "the same 16 positions: C feature oracle (1 thread) %.2f ms + torch fp32 3x32 forward %.2f ms" % (1e3 * t_port, 1e3 * t_fwd)}
        if ref:
            f = ref["single"]["value"]
            line["cpu_baseline"]["reference"] = {"value": 1.0 / (1.0 / f + t_fwd / n), "unit": "positions/s", "cores": 1, "kind": "reference",
                                                 "features_only": f, "planes_equal_c_port": ref["single"]["planes_equal_c_port"],
                                                 "sample": "Preprocess.state_to_tensor of the unmodified reference on the same 16 positions ("
                                                           + ref["single"]["sample"] + ") + the torch fp32 forward"}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=4096, help="positions per GPU per step")
    ap.add_argument("--pool", type=int, default=4, help="distinct input batches cycled through")
    ap.add_argument("--dtype", default="fp16c32", choices=["fp16c32", "fp16c64", "fp16", "bf16", "fp16x3"],
                    help="operand precision of the convolutions: fp16c32 (default; fp16 with the 32 most energetic input channels of "
                         "layers 2.. corrected by their fp16 remainders -- holds the 99.5 %% top-1 gate), fp16 / bf16 (plain 16-bit "
                         "operands, 1.3x faster, at / below the gate on random-init weights), fp16x3 (fp32-like, 3x the MMA work)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=8192, help="positions the CPU port is timed on (~15 s on 16 cores)")
    ap.add_argument("--ref-sample", type=int, default=4096, help="positions per step of the --impl reference arm (the GPU arm's batch)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="policy", choices=["policy", "features", "policy+value", "value", "cfg1"],
                    help="policy: the flagship (features + 13x192 policy forward); features: BASELINE config 4; "
                         "policy+value: BASELINE config 5 (adds the 49-plane 13x192 value forward to every step)")
    ap.add_argument("--serial", action="store_true", help="also time the single-stream rag_eval_batch step")
    ap.add_argument("--conv-pair", type=int, default=0, help="0: one CTA per tile (default); 1: convolutions on CTA pairs (cta_group::2; faster, but "
                                                               "sporadically stalls beside the pipeline's ladder kernels -- see csrc/conv.cu)")
    ap.add_argument("--no-pipeline", action="store_true", help="debug: skip the pipelined passes")
    ap.add_argument("--e2e-graphs", type=int, default=1, help="1: the end-to-end pass replays each pipeline stage as a CUDA graph")
    ap.add_argument("--slots", type=int, default=None, help="batches in flight in the pipeline (default: 2, or 4 below 2048 positions)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "features":
        if args.batch == 4096:
            args.batch = 32768
        if args.steps == 100:
            args.steps = 32                 # 32 x 32,768 = 1,048,576 distinct positions per GPU
        run_features(args)
    elif args.workload == "cfg1":
        run_cfg1(args)
    elif args.workload == "value":
        if args.batch == 4096:
            args.batch = 1024
        run_gpu(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
