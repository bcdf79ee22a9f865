"""Timeline of the pipeline: when does each stage of each batch start/finish on the GPU?
usage: pipeline_timeline.py [batch] [slots] [helper CTAs]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState, _cabi
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.pipeline import EvalPipeline
from rocalphago_b200.preprocessing import DEFAULT_FEATURES

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
if len(sys.argv) > 3:
    _cabi.lib().rag_set_ladder_helpers(int(sys.argv[3]))
net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype=os.environ.get("DTYPE", "fp16c32"), seed=1)
batches = [BatchedGameState.random_playouts(n, seed=5 + i) for i in range(4)]
pipe = EvalPipeline(net, n, slots=int(sys.argv[2]) if len(sys.argv) > 2 else None)
for i in range(8):
    pipe.submit(batches[i % 4].packed)
pipe.drain()
t0 = torch.cuda.Event(enable_timing=True); t0.record(); torch.cuda.synchronize()
pipe.trace = []
for k in range(16):
    pipe.submit(batches[k % 4].packed)
pipe.drain(); torch.cuda.synchronize()
print("batch | features start | analysis end | ladders+emit end | forward start .. end   (ms since t0)")
prev = None
for k, tr in enumerate(pipe.trace):
    t = [t0.elapsed_time(e) for e in tr]
    print("%5d | %7.2f | %7.2f | %7.2f | %7.2f .. %7.2f  (forward %.2f, step %.2f)" % (
        k, t[0], t[1], t[2], t[3], t[4], t[4] - t[3], (t[4] - prev) if prev else 0))
    prev = t[4]
