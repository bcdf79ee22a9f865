"""Timeline of the two-stage pipeline: when does each stage of each batch start/finish on the GPU?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.pipeline import EvalPipeline
from rocalphago_b200.preprocessing import DEFAULT_FEATURES

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype="fp16", seed=1)
batches = [BatchedGameState.random_playouts(n, seed=20261106 + 7919 * i) for i in range(4)]
pipe = EvalPipeline(net, n)
for i in range(4):
    pipe.submit(batches[i % 4].packed)
pipe.drain()
# instrument: wrap the events with timing-enabled twins
K = 8
marks = []
E = lambda: torch.cuda.Event(enable_timing=True)
t0 = E(); t0.record(); torch.cuda.synchronize()
orig_submit = pipe.submit
for k in range(K):
    slot = pipe.slots[pipe.count % 2]
    a0, a1, f1, n0, n1 = E(), E(), E(), E(), E()
    # start markers: recorded on the stage's stream after its waits would be ideal; approximate by recording
    # "done" events right after the stage (the pipeline's own events) and "start" right before enqueue
    with torch.cuda.stream(pipe.s_ana):
        if slot.used:
            pipe.s_ana.wait_event(slot.net_done)
        a0.record(pipe.s_ana)
    t = pipe.submit(batches[k % 4].packed)
    a1.record(pipe.s_ana); f1.record(pipe.s_feat); n1.record(pipe.s_net)
    marks.append((a0, a1, f1, n1))
pipe.drain(); torch.cuda.synchronize()
print("batch | analysis start..end | ladder+emit end | net end   (ms since t0)")
prev_n = None
for k, (a0, a1, f1, n1) in enumerate(marks):
    print("%5d | %7.2f .. %7.2f | %7.2f | %7.2f   (step %.2f)" % (
        k, t0.elapsed_time(a0), t0.elapsed_time(a1), t0.elapsed_time(f1), t0.elapsed_time(n1),
        (t0.elapsed_time(n1) - prev_n) if prev_n else 0))
    prev_n = t0.elapsed_time(n1)
