"""analyze+ladder time vs the hand-over threshold, on a few batches."""
import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState, Preprocess, _cabi
from rocalphago_b200.state import _ptr, _stream
L = _cabi.lib()
pp = Preprocess()
for n, seeds in ((256, (7, 20261106, 3)), (4096, (7, 20261106, 20261106 + 7919, 20261106 + 2 * 7919)), (32768, (7,))):
    batches = [BatchedGameState.random_playouts(n, seed=s) for s in seeds]
    ws = pp._ws.get(n, batches[0].device)
    for thr in (-1, 32, 48, 96, 192, 384):
        L.rag_set_ladder_handover(thr)
        res = []
        for b in batches:
            f = lambda: _cabi.check(L.rag_features_analyze(_ptr(b.packed), n, 19, 1, _ptr(ws), ws.numel(), _stream()))
            for _ in range(2): f()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): f()
            e1.record(); torch.cuda.synchronize()
            res.append(e0.elapsed_time(e1) / 5)
        print("n=%5d handover=%4d  ms per batch: %s" % (n, thr, "  ".join("%.3f" % r for r in res)), flush=True)
