"""How long do ladder searches get?  Largest number of moves tried along one line of a search, per batch of
device-generated random-playout positions (the bench's seeds), with the time each batch's ladder stage takes.
usage: ladder_longest.py [first rank] [last rank] [batches per rank] [handover]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rocalphago_b200 import BatchedGameState, _cabi
from rocalphago_b200.state import _ptr, _stream
L = _cabi.lib()
r0, r1 = int(sys.argv[1]), int(sys.argv[2])
per = int(sys.argv[3]) if len(sys.argv) > 3 else 32
if len(sys.argv) > 4:
    L.rag_set_ladder_handover(int(sys.argv[4]))
n = 32768
ws = torch.empty(int(L.rag_features_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
longest, times = [], []
for rank in range(r0, r1 + 1):
    for i in range(per):
        b = BatchedGameState.random_playouts(n, seed=20261106 + 7919 * (rank * 64 + i))
        torch.cuda.synchronize()
        t0 = time.time()
        _cabi.check(L.rag_features_analyze(_ptr(b.packed), n, 19, 1, _ptr(ws), ws.numel(), _stream()))
        torch.cuda.synchronize()
        times.append(time.time() - t0)
        longest.append(L.rag_features_ladder_longest(_ptr(ws), _stream()))
        cut = L.rag_features_ladder_cutoffs(_ptr(ws), _stream())
        if cut or times[-1] > 0.03:
            print("rank %d batch %d: longest %d, cut off %d, %.3f s" % (rank, i, longest[-1], cut, times[-1]), flush=True)
longest = np.array(longest)
print("batches %d (%d positions): longest line per batch: median %d, p90 %d, p99 %d, max %d; sorted top 10: %s; batch time median %.1f ms max %.1f ms" % (
    len(longest), len(longest) * n, np.median(longest), np.percentile(longest, 90), np.percentile(longest, 99), longest.max(),
    np.sort(longest)[-10:].tolist(), 1e3 * np.median(times), 1e3 * max(times)))
