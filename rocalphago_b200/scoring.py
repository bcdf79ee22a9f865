"""Winners and rollouts on the device (GameState.get_winner, go.py:482-506; MCTS._evaluate_rollout,
mcts.py:171-189 with uniformly random sensible moves)."""
import ctypes

import numpy as np
import torch

from . import _cabi
from .state import _ptr, _require_cuda, _stream


def random_rollout_results(batch, limit=500, seed=0, komi=7.5):
    """numpy float64 (n,): +1 if the player to move in batch[i] wins a random playout from there
    (`limit` moves at most), -1 if the opponent does, 0 on a tie.  `batch` is not modified."""
    _require_cuda()
    n = len(batch)
    out = torch.zeros(n, dtype=torch.int32, device=batch.device)
    with torch.cuda.device(batch.device):
        _cabi.check(_cabi.lib().rag_rollout_winners(_ptr(batch.packed), n, batch.size, int(limit), ctypes.c_uint64(int(seed)),
                                                    ctypes.c_float(komi), _ptr(out), _stream()))
    return out.cpu().numpy().astype(np.float64)


def winners(batch, komi=7.5):
    """get_winner() of every position (no passes counted): +1 BLACK, -1 WHITE, 0 tie."""
    rel = random_rollout_results(batch, limit=0, komi=komi)
    player = batch.packed[:, 5 * 32 + 1].cpu().numpy()
    return (rel * player).astype(np.int32)
