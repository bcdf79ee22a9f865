import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import cnn_oracle, go_oracle
from rocalphago_b200.models import CNNPolicy, CNNValue
feats = go_oracle.DEFAULT_FEATURES + ["color"]
b, a, ko, pl = go_oracle.random_playouts(32, seed=11, max_ply=361, stride=5)
planes = go_oracle.features(b, a, ko, pl, feats)
for dtype in ("bf16", "fp16"):
    net = CNNValue(feats, dtype=dtype, seed=3)
    w = net.model.get_weights()
    got = net.forward(planes.astype(np.float32)).reshape(-1)
    lg = net.model.last_logits(32).cpu().numpy()
    # oracle logits: policy-style forward of the conv part
    wl = cnn_oracle.forward(w[:-4] + [np.zeros(361, np.float32)], planes, net.model.specs, head="policy", return_logits=True)
    want = cnn_oracle.forward(w, planes, net.model.specs, head="value")
    print(dtype, "logits: max abs err %.3g, oracle mean %.3g std %.3g, device mean %.3g" % (np.abs(lg - wl).max(), wl.mean(), wl.std(), lg.mean()))
    d1 = lg @ w[-4] + w[-3]; d2 = d1 @ w[-2] + w[-1]
    print(dtype, "value from device logits on cpu: max err vs oracle %.3g ; device value err %.3g" % (np.abs(np.tanh(d2).reshape(-1) - want).max(), np.abs(got - want).max()))
    print(got[:6], want[:6])
