"""One stand-alone analysis + ladder pass over 4,096 random-playout positions (for ncu captures)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState, Preprocess, _cabi
from rocalphago_b200.state import _ptr, _stream
L = _cabi.lib(); pp = Preprocess()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
b = BatchedGameState.random_playouts(n, seed=7)
ws = pp._ws.get(n, b.device)
for _ in range(3):
    _cabi.check(L.rag_features_analyze(_ptr(b.packed), n, 19, 1, _ptr(ws), ws.numel(), _stream()))
torch.cuda.synchronize()
print("ok")
