import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState, _cabi
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.pipeline import EvalPipeline
from rocalphago_b200.preprocessing import DEFAULT_FEATURES
L = _cabi.lib()
net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype="fp16", seed=1)
for n in (64, 256, 1024):
    bs = [BatchedGameState.random_playouts(n, seed=5 + i) for i in range(4)]
    for helpers in (32, 8):
        L.rag_set_ladder_helpers(helpers)
        for slots in (4, 6, 8, 12):
            pipe = EvalPipeline(net, n, graphs=True, slots=slots)
            for i in range(16): pipe.submit(bs[i % 4].packed)
            pipe.drain(); torch.cuda.synchronize()
            K = 200
            t0 = time.time()
            for i in range(K): pipe.submit(bs[i % 4].packed)
            pipe.drain(); torch.cuda.synchronize()
            t_all = time.time() - t0
            print("carveout=%s n=%5d helpers=%4d slots=%d  %.1f us/step -> %.0f pos/s" % (os.environ.get("RAG_CARVEOUT"), n, helpers, slots, t_all / K * 1e6, n * K / t_all))
