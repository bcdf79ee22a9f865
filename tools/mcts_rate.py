"""Playouts per second of BatchedMCTS with the 13x192 policy and value networks on a 19x19 middle-game
position, for several wave sizes (reference defaults: playout_depth 20, lmbda 0 here)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rocalphago_b200 import BatchedGameState
from rocalphago_b200.mcts import BatchedMCTS
from rocalphago_b200.models import CNNPolicy, CNNValue
from rocalphago_b200.preprocessing import DEFAULT_FEATURES

pol = CNNPolicy(DEFAULT_FEATURES, layers=12, filters_per_layer=192, seed=1)
val = CNNValue(DEFAULT_FEATURES + ["color"], layers=12, filters_per_layer=192, seed=2)
pos = BatchedGameState.random_playouts(64, seed=11, max_ply=120, stride=8)
root = BatchedGameState(pos.packed[40:41].clone(), 19)
for batch, n in ((1, 16), (16, 128), (64, 512), (256, 2048), (1024, 8192)):
    s = BatchedMCTS(pol, val, playout_depth=20, n_playout=n, batch=batch)
    s.get_move(root); torch.cuda.synchronize()                     # warm-up (also grows the tree: a second search is the steady state)
    s = BatchedMCTS(pol, val, playout_depth=20, n_playout=n, batch=batch)
    t0 = time.perf_counter(); mv = s.get_move(root); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("batch %4d: %5d playouts in %6.2f s = %7.1f playouts/s, %6d policy evaluations, move %s" % (batch, n, dt, n / dt, s.evaluations, mv))

from rocalphago_b200.mcts import ParallelMCTS
roots = BatchedGameState(pos.packed[16:48].clone(), 19)
for games, batch, n in ((16, 64, 512), (32, 128, 1024)):
    sub = BatchedGameState(roots.packed[:games].clone(), 19)
    par = ParallelMCTS(pol, val, games=games, playout_depth=20, n_playout=n, batch=batch)
    par.get_moves(sub); torch.cuda.synchronize()
    par = ParallelMCTS(pol, val, games=games, playout_depth=20, n_playout=n, batch=batch)
    t0 = time.perf_counter(); mv = par.get_moves(sub); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    ev = sum(s.evaluations for s in par.searches)
    print("parallel %2d trees x batch %3d: %6d playouts in %6.2f s = %8.1f playouts/s, %7d policy evaluations (%.0f /s)" % (
        games, batch, games * n, dt, games * n / dt, ev, ev / dt))
