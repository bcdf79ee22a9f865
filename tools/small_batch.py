"""Small-batch throughput of the pipeline vs the number of slots and of ladder helper CTAs; plus the
serial per-stage device times of the same batches."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState, _cabi
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.pipeline import EvalPipeline
from rocalphago_b200.preprocessing import DEFAULT_FEATURES
from rocalphago_b200.state import _ptr, _stream
L = _cabi.lib()
E = lambda: torch.cuda.Event(enable_timing=True)
net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype="fp16", seed=1)
for n in (64, 256, 1024):
    bs = [BatchedGameState.random_playouts(n, seed=5 + i) for i in range(4)]
    pipe = EvalPipeline(net, n, slots=2)
    sl = pipe.slots[0]
    for b in bs:
        sl.packed.copy_(b.packed)
        ts = []
        for fn in (EvalPipeline._stage_analysis, EvalPipeline._stage_ladders_emit, EvalPipeline._stage_forward):
            for _ in range(2): fn(pipe, sl, torch.cuda.current_stream())
            torch.cuda.synchronize(); e0, e1 = E(), E(); e0.record()
            for _ in range(5): fn(pipe, sl, torch.cuda.current_stream())
            e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) / 5)
        print("n=%d serial: analysis %.3f  ladders+emit %.3f  forward %.3f ms" % (n, *ts))
    for helpers in (-1, 148, 32):
        L.rag_set_ladder_helpers(helpers)
        for slots in (2, 4):
            pipe = EvalPipeline(net, n, graphs=True, slots=slots)
            for i in range(16): pipe.submit(bs[i % 4].packed)
            pipe.drain(); torch.cuda.synchronize()
            K = 200
            t0 = time.time()
            for i in range(K): pipe.submit(bs[i % 4].packed)
            pipe.drain(); torch.cuda.synchronize()
            t_all = time.time() - t0
            print("n=%5d helpers=%4d slots=%d  %.1f us/step -> %.0f pos/s" % (n, helpers, slots, t_all / K * 1e6, n * K / t_all))
    L.rag_set_ladder_helpers(-1)
