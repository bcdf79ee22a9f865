"""How close is the uint8 emit kernel to what HBM allows?  Times it alone (256 MiB written before every launch)
beside a plain fill and a plain copy of the same number of bytes."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rocalphago_b200 import BatchedGameState, Preprocess, _cabi
from rocalphago_b200 import preprocessing as P
from rocalphago_b200.state import _ptr, _stream
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
L = _cabi.lib(); pp = Preprocess()
b = BatchedGameState.random_playouts(n, seed=5)
ws = torch.empty(int(L.rag_features_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
out = torch.empty((n, 48, 19, 19), dtype=torch.uint8, device="cuda")
src = np.ascontiguousarray(pp.plane_src, np.int32); vp = ctypes.c_void_p
_cabi.check(L.rag_features_analyze(_ptr(b.packed), n, 19, 1, _ptr(ws), ws.numel(), _stream()))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timed(fn, reps=7):
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); c.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(c))
    return float(np.median(ts))
emit = timed(lambda: _cabi.check(L.rag_features_emit(_ptr(ws), n, 19, src.ctypes.data_as(vp), len(src), P.LAYOUT_NCHW_U8, 0, 0, _ptr(out), _stream())))
fill = timed(lambda: out.zero_())
other = torch.empty_like(out)
copy = timed(lambda: other.copy_(out))
B = n * 17328
print("variant=%s n=%d: emit %.4f ms (%.0f GB/s written, %.0f GB/s algorithmic 17,568 B/pos) | fill %.4f ms (%.0f GB/s) | copy %.4f ms (%.0f GB/s r+w)" % (
    os.environ.get("RAG_EMIT_VARIANT", "0"), n, emit, B / emit / 1e6, n * 17568 / emit / 1e6, fill, B / fill / 1e6, copy, 2 * B / copy / 1e6))
