"""Ad-hoc device timing of the feature kernels (CUDA events), per stage."""
import sys
import os
import ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState, Preprocess, _cabi
from rocalphago_b200 import preprocessing as P
from rocalphago_b200.state import _ptr, _stream


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    for n in (256, 4096, 32768):
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        batch = BatchedGameState.random_playouts(n, seed=7)
        t1.record(); torch.cuda.synchronize()
        print("n=%d playouts generated in %.1f ms" % (n, t0.elapsed_time(t1)))
        pp = Preprocess()
        L = _cabi.lib()
        ws = pp._ws.get(n, batch.device)
        noladder = lambda: _cabi.check(L.rag_features_analyze(_ptr(batch.packed), n, 19, 0, _ptr(ws), ws.numel(), _stream()))
        full = lambda: _cabi.check(L.rag_features_analyze(_ptr(batch.packed), n, 19, 1, _ptr(ws), ws.numel(), _stream()))
        out_u8 = torch.empty((n, 48, 19, 19), dtype=torch.uint8, device="cuda")
        out_cv = pp.emit(ws, n, 19, P.LAYOUT_CONV_IN)
        emit_u8 = lambda: pp.emit(ws, n, 19, P.LAYOUT_NCHW_U8, out=out_u8)
        emit_cv = lambda: pp.emit(ws, n, 19, P.LAYOUT_CONV_IN, out=out_cv)
        a = timeit(noladder); b = timeit(full); c = timeit(emit_u8); d = timeit(emit_cv)
        items = int(ws[:8].view(torch.int32)[0].item())
        print("n=%d analyze %.3f ms (%.0f pos/s) | analyze+ladder %.3f ms (%.0f pos/s, %d items) | emit u8 %.3f ms (%.1f GB/s) | emit conv %.3f ms (%.1f GB/s)" % (
            n, a, n / a * 1e3, b, n / b * 1e3, items, c, n * 48 * 361 / c / 1e6, d, n * 361 * 128 / d / 1e6))


if __name__ == "__main__":
    main()
