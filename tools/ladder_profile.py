"""Per-item cost of the ladder reader: builds a timing variant of the library
(-DRAG_LADDER_TIMING: every queue slot is overwritten with the item's clock64 cycles / 16)."""
import ctypes
import os
import subprocess
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "tools", "_build", "libfeatures_timing.so")


def build():
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
                           "-DRAG_LADDER_TIMING", "-shared", "-Xcompiler", "-fPIC", "-o", SO,
                           os.path.join(ROOT, "rocalphago_b200", "csrc", "features.cu")])


def main():
    from rocalphago_b200 import BatchedGameState
    L = ctypes.CDLL(SO)
    L.rag_features_workspace_bytes.restype = ctypes.c_size_t
    for n in (256, 4096):
        batch = BatchedGameState.random_playouts(n, seed=20261106)
        nbytes = L.rag_features_workspace_bytes(n)
        ws = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
        vp = ctypes.c_void_p
        for rep in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = L.rag_features_analyze(vp(batch.packed.data_ptr()), n, 19, 1, vp(ws.data_ptr()), ctypes.c_size_t(nbytes), None)
            e1.record()
            torch.cuda.synchronize()
            assert rc == 0
        items = int(ws[:4].view(torch.int32)[0])
        q = ws[256 + n * 50 * 32 * 4: 256 + n * 50 * 32 * 4 + items * 4].view(torch.int32).cpu().numpy().astype(np.int64) * 16
        print("n=%d items=%d  analyze+ladder %.3f ms" % (n, items, e0.elapsed_time(e1)))
        print("  cycles/item: mean %.0f  median %.0f  p90 %.0f  p99 %.0f  p99.9 %.0f  max %.0f   sum %.3g" % (
            q.mean(), np.median(q), np.percentile(q, 90), np.percentile(q, 99), np.percentile(q, 99.9), q.max(), q.sum()))
        print("  top 8 items (Mcycles):", np.round(np.sort(q)[-8:] / 1e6, 3).tolist())
        print("  ideal if perfectly spread over 4736 warps: %.3f ms at 1.9 GHz; longest single item %.3f ms" % (
            q.sum() / 4736 / 1.9e6, q.max() / 1.9e6))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "build":
        build()
    else:
        main()
