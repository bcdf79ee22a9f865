"""Batched policy players on the device (drop-in for the batched half of AlphaGo/ai.py).

`ProbabilisticPolicyPlayer.get_moves(states)` (ai.py:107-133) is the reference's only real
batching; here the whole loop -- features, policy forward, move selection, do_move -- stays on
the GPU, which is what a self-play generator (reinforcement_policy_trainer.py:21-86's
run_n_games, minus the training) needs.
"""
import ctypes

import numpy as np
import torch

from . import _cabi
from .state import BatchedGameState, _ptr, _stream

PASS_MOVE = None


class BatchedPolicyPlayer(object):
    """Greedy or temperature-sampling player over a BatchedGameState.

    temperature, greedy_start, move_limit follow ProbabilisticPolicyPlayer (ai.py:57-66); pass
    greedy=True for GreedyPolicyPlayer's behaviour.  Moves are flat indices x*S+y, -1 = pass.
    """

    def __init__(self, policy, temperature=1.0, greedy=False, greedy_start=None, move_limit=None, seed=0):
        assert temperature > 0.0
        self.policy, self.temperature, self.greedy = policy, float(temperature), bool(greedy)
        self.greedy_start, self.move_limit, self.seed = greedy_start, move_limit, int(seed)
        self.calls = 0

    def get_moves(self, batch, ply=0):
        """int32 CUDA tensor (n,) of moves for every position of `batch` (all at ply `ply`)."""
        if not isinstance(batch, BatchedGameState):
            batch = BatchedGameState.from_states(batch)
        n = len(batch)
        if self.move_limit is not None and ply > self.move_limit:
            return torch.full((n,), -1, dtype=torch.int32, device=batch.device)
        model = self.policy.model
        probs = model.eval_batch_device(batch, self.policy.preprocessor.plane_src)
        ws = model.features_workspace()
        moves = torch.empty(n, dtype=torch.int32, device=batch.device)
        greedy = self.greedy or (self.greedy_start is not None and ply >= self.greedy_start)
        self.calls += 1
        with torch.cuda.device(batch.device):
            _cabi.check(_cabi.lib().rag_select_moves(_ptr(probs), _ptr(ws), n, batch.size, ctypes.c_float(self.temperature),
                                                     int(greedy), ctypes.c_uint64(self.seed * 1000003 + self.calls),
                                                     _ptr(moves), _stream()))
        return moves

    def get_moves_list(self, states):
        """The reference's return type: a list of (x, y) tuples / None per state."""
        batch = states if isinstance(states, BatchedGameState) else BatchedGameState.from_states(states)
        mv = self.get_moves(batch).cpu().numpy()
        S = batch.size
        return [PASS_MOVE if m < 0 else (int(m) // S, int(m) % S) for m in mv]


def play_games(player, n_games, size=19, max_moves=500, opponent=None, device=None, record=False):
    """n_games simultaneous games from the empty board, both sides played by `player` (or black by
    `player`, white by `opponent`), entirely on the device.  A game stops moving once both sides
    have passed in a row; the loop ends when all have, or after max_moves plies.
    Returns the final BatchedGameState (and, with record=True, the int32 (plies, n_games) move log)."""
    dev = torch.device(device or "cuda")
    packed = torch.zeros((n_games, 192), dtype=torch.int32, device=dev)
    packed[:, 160] = -1          # ko: none
    packed[:, 161] = 1           # black to move
    packed[:, 162] = size
    batch = BatchedGameState(packed, size)
    last_pass = torch.zeros(n_games, dtype=torch.bool, device=dev)
    over = torch.zeros(n_games, dtype=torch.bool, device=dev)
    log = []
    for ply in range(max_moves):
        who = player if (opponent is None or ply % 2 == 0) else opponent
        moves = who.get_moves(batch, ply=ply)
        moves = torch.where(over, torch.full_like(moves, -1), moves)
        status = batch.do_move(moves)
        assert int((status != 0).sum()) == 0, "a selected move was refused"
        is_pass = moves < 0
        over = over | (is_pass & last_pass)
        last_pass = is_pass
        if record:
            log.append(moves.clone())
        if bool(over.all()):
            break
    return (batch, torch.stack(log)) if record else batch
