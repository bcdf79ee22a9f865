"""Smallest end-to-end case for compute-sanitizer: features (with ladders) + a 3-layer policy on 48 positions."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from rocalphago_b200 import BatchedGameState, Preprocess
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.preprocessing import DEFAULT_FEATURES
batch = BatchedGameState.random_playouts(48, seed=3, max_ply=300, stride=6)
planes = Preprocess().to_device(batch).cpu().numpy()
net = CNNPolicy(DEFAULT_FEATURES, layers=3, filters_per_layer=64, dtype="fp16", seed=1)
probs = net.model.eval_batch_device(batch, net.preprocessor.plane_src).cpu().numpy()
st = batch.do_move(torch.full((48,), -1, dtype=torch.int32)).cpu().numpy()
torch.cuda.synchronize()
print("ok", planes.sum(), float(probs.sum()), int(st.sum()))
