"""How long does the host take to enqueue one pipelined step, vs the GPU time of the step?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.pipeline import EvalPipeline
from rocalphago_b200.preprocessing import DEFAULT_FEATURES
for n, graphs, slots in [(n, g, sl) for n in (64, 256, 1024) for g in (False, True) for sl in (2, 4, 6)]:
    net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype="fp16", seed=1)
    bs = [BatchedGameState.random_playouts(n, seed=5 + i) for i in range(4)]
    pipe = EvalPipeline(net, n, graphs=graphs, slots=slots)
    for i in range(16): pipe.submit(bs[i % 4].packed)
    pipe.drain(); torch.cuda.synchronize()
    K = 200
    t0 = time.time()
    for i in range(K): pipe.submit(bs[i % 4].packed)
    t_host = time.time() - t0
    pipe.drain(); torch.cuda.synchronize()
    t_all = time.time() - t0
    print("graphs=%d slots=%d " % (graphs, slots) + "n=%5d  host enqueue %.1f us/step   wall incl. GPU %.1f us/step   -> %.0f pos/s" % (n, t_host / K * 1e6, t_all / K * 1e6, n * K / t_all))
