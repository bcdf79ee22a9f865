"""Which part of the feature stage slows the forward down when they overlap?"""
import os, sys, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState, _cabi
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200 import pipeline as PL
from rocalphago_b200.pipeline import EvalPipeline, _ptr, LAYOUT_CONV_IN
from rocalphago_b200.preprocessing import DEFAULT_FEATURES
L = _cabi.lib()
L.rag_set_ladder_helpers(8)
net = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype="fp16", seed=1)
orig_feat, orig_ana = EvalPipeline._stage_ladders_emit, EvalPipeline._stage_analysis
def emit_only(self, slot, stream):
    vp = ctypes.c_void_p; st = vp(stream.cuda_stream)
    _cabi.check(L.rag_features_emit(_ptr(slot.fws), self.n, self.size, self.plane_src.ctypes.data_as(vp), len(self.plane_src),
                                    LAYOUT_CONV_IN, self.dtype_code, self.channels, _ptr(slot.rows), st))
def nothing(self, slot, stream): pass
def run(tag, n, slots=int(os.environ.get('SLOTS', '4'))):
    bs = [BatchedGameState.random_playouts(n, seed=5 + i) for i in range(4)]
    pipe = EvalPipeline(net, n, graphs=False, slots=slots)
    for i in range(16): pipe.submit(bs[i % 4].packed)
    pipe.drain(); torch.cuda.synchronize()
    K = int(os.environ.get('K', '200'))
    pipe.trace = []
    t0 = time.time()
    for i in range(K): pipe.submit(bs[i % 4].packed)
    pipe.drain(); torch.cuda.synchronize()
    t_all = time.time() - t0
    fw = sorted(tr[3].elapsed_time(tr[4]) for tr in pipe.trace)[K // 2]
    ft = sorted(tr[0].elapsed_time(tr[2]) for tr in pipe.trace)[K // 2]
    print("%-28s n=%5d  %.1f us/step  median forward %.3f ms  median feature stage %.3f ms" % (tag, n, t_all / K * 1e6, fw, ft))
for n in [int(x) for x in os.environ.get('NS', '64,256,1024').split(',')]:
    run("full", n)
    if os.environ.get("ONLY_FULL"): continue
    L.rag_set_ladder_handover(-1); run("no hand-over", n); L.rag_set_ladder_handover(96)
    EvalPipeline._stage_ladders_emit = emit_only; run("no ladder kernel", n)
    EvalPipeline._stage_ladders_emit = nothing; run("analysis only", n)
    EvalPipeline._stage_analysis = nothing; run("forward only", n)
    EvalPipeline._stage_ladders_emit, EvalPipeline._stage_analysis = orig_feat, orig_ana
