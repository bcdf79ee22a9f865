"""What bounds the convolution kernel?  Sustained time of the 13x192 forward (convolutions only, from pixel
rows already in HBM) under the timing knobs of csrc/conv.cu:

    RAG_CONV_DEBUG   1 skip weight loads   2 skip activation loads   3 both   8 skip the epilogue's stores
    RAG_CONV_STAGES  fewer ring stages
    python tools/conv_bound.py [dtype] [pair] [n] [forwards]

Prints mean ms per 3x3 layer over the second half of `forwards` back-to-back forwards (the part is at its
power cap by then) and the SM clock nvidia-smi reports meanwhile.  Results under RAG_CONV_DEBUG are garbage
by design -- only the time is read."""
import ctypes, os, subprocess, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rocalphago_b200 import _cabi
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.preprocessing import DEFAULT_FEATURES
from rocalphago_b200.state import _ptr, _stream

dtype = sys.argv[1] if len(sys.argv) > 1 else "fp16"
pair = int(sys.argv[2]) if len(sys.argv) > 2 else 0
n = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 60
L = _cabi.lib()
net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype=dtype, seed=3)
dev = torch.device("cuda", 0)
h = net.model._ensure_handle(dev)
L.rag_set_conv_pair(pair)
rows, ch = int(L.rag_net_input_rows(h, n)), int(L.rag_net_input_channels(h))
x = (torch.rand((rows, ch), device=dev) < 0.2).to(torch.float16)
out = torch.empty((n, 361), dtype=torch.float32, device=dev)
ws = torch.empty(int(L.rag_net_workspace_bytes(h, n)), dtype=torch.uint8, device=dev)
_cabi.check(L.rag_net_set_profiling(h, 1))
clk = []
def sample():
    p = subprocess.Popen(["nvidia-smi", "-i", "0", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "100"],
                         stdout=subprocess.PIPE, text=True)
    sample.p = p
    for line in p.stdout:
        try:
            a, b = line.split(","); clk.append((float(a), float(b)))
        except Exception:
            pass
threading.Thread(target=sample, daemon=True).start()
time.sleep(0.3)
per = []
t0 = time.time()
for r in range(reps):
    _cabi.check(L.rag_net_forward_rows(h, _ptr(x), n, _ptr(out), _ptr(ws), ws.numel(), _stream()))
    if r >= reps // 2:
        ms = (ctypes.c_float * 12)()
        _cabi.check(L.rag_net_layer_times(h, ms, 12))
        per.append(list(ms))
torch.cuda.synchronize()
wall = time.time() - t0
sample.p.terminate()
per = np.asarray(per)
busy = [c for c in clk if c[1] > 500]
print("dtype=%s pair=%d n=%d debug=%s stages=%s : 3x3 layer %.4f ms  layer1 %.4f ms  all %.3f ms | sm %.0f MHz %.0f W (%d samples)" % (
    dtype, pair, n, os.environ.get("RAG_CONV_DEBUG", "0"), os.environ.get("RAG_CONV_STAGES", "-"),
    per[:, 1:].mean(), per[:, 0].mean(), per.sum(1).mean(),
    np.median([c[0] for c in busy]) if busy else -1, np.median([c[1] for c in busy]) if busy else -1, len(busy)), flush=True)
