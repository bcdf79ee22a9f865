"""Latency of the reference-facing calls (host objects in, Python lists out) at small sizes:
BASELINE config 1 (16 positions, policy layers=3 filters=32) and single-state eval_state."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rocalphago_b200 import BatchedGameState
from rocalphago_b200.models import CNNPolicy, CNNValue
from rocalphago_b200.preprocessing import DEFAULT_FEATURES, Preprocess

def timeit(fn, reps=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3

views = BatchedGameState.random_playouts(64, seed=3).views()
small = CNNPolicy(DEFAULT_FEATURES, layers=3, filters_per_layer=32, seed=1)
big = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, seed=1)
val = CNNValue(DEFAULT_FEATURES + ["color"], layers=12, filters_per_layer=192, seed=2)
pp = Preprocess()
print("config 1: batch_eval_state(16 states), 3x32 policy : %.3f ms" % timeit(lambda: small.batch_eval_state(views[:16])))
print("batch_eval_state(16 states), 13x192 policy          : %.3f ms" % timeit(lambda: big.batch_eval_state(views[:16])))
print("eval_state(1 state), 13x192 policy                  : %.3f ms" % timeit(lambda: big.eval_state(views[5])))
print("eval_state(1 state), 13x192 value                   : %.3f ms" % timeit(lambda: val.eval_state(views[5])))
print("Preprocess.state_to_tensor(1 state)                 : %.3f ms" % timeit(lambda: pp.state_to_tensor(views[5])))
print("batch_eval_state(64 states), 13x192 policy          : %.3f ms" % timeit(lambda: big.batch_eval_state(views)))
