"""Policy players under the reference's own names (drop-in for AlphaGo/ai.py).

`GreedyPolicyPlayer`, `ProbabilisticPolicyPlayer` and `MCTSPlayer` take the constructor arguments of
ai.py:8-16, 41-55, 136-140 and answer `get_move(state)` / `get_moves(states)` with (x, y) tuples or
PASS_MOVE (None), so callers written against the reference (interface/gtp_wrapper.py,
training/reinforcement_policy_trainer.py) need no change.  Underneath, the list of sensible moves, the
policy forward, the temperature and the choice all run on the device (BatchedPolicyPlayer / rag_select_moves);
states are GameState-like objects (`.size .board .stone_ages .ko .current_player`, and `.history` where
the move-limit / pass-when-offered rules read it) or a BatchedGameState.

Differences a caller can see: the random draw comes from a seeded device generator, not numpy's global
one (same distribution, other numbers), and `policy_function` / `value_function` must be this package's
CNNPolicy / CNNValue objects (or their `eval_state` bound methods) -- the kernels cannot call back into
arbitrary Python.
"""
import numpy as np

from .mcts import MCTS
from .players import BatchedPolicyPlayer
from .state import BatchedGameState

PASS_MOVE = None          # go.py:8


def _plies(state):
    return len(getattr(state, "history", ()))


def _offered_pass(state):
    h = getattr(state, "history", ())
    return len(h) > 100 and h[-1] is PASS_MOVE


class GreedyPolicyPlayer(object):
    """A player that uses a greedy policy (i.e. chooses the highest probability move each turn) (ai.py:8-38)"""

    def __init__(self, policy_function, pass_when_offered=False, move_limit=None):
        self.policy = policy_function
        self.pass_when_offered = pass_when_offered
        self.move_limit = move_limit
        self._player = BatchedPolicyPlayer(policy_function, greedy=True)

    def get_move(self, state):
        if self.move_limit is not None and _plies(state) > self.move_limit:
            return PASS_MOVE
        if self.pass_when_offered and _offered_pass(state):
            return PASS_MOVE
        return self._player.get_moves_list([state])[0]


class ProbabilisticPolicyPlayer(object):
    """A player that samples a move in proportion to the probability given by the policy; `temperature`
    pushes it towards random (high) or greedy (low) play (ai.py:41-133)."""

    def __init__(self, policy_function, temperature=1.0, pass_when_offered=False, move_limit=None, greedy_start=None,
                 seed=0):
        assert(temperature > 0.0)
        self.policy = policy_function
        self.move_limit = move_limit
        self.beta = 1.0 / temperature
        self.pass_when_offered = pass_when_offered
        self.greedy_start = greedy_start
        self._sample = BatchedPolicyPlayer(policy_function, temperature=temperature, seed=seed)
        self._greedy = BatchedPolicyPlayer(policy_function, greedy=True)

    def apply_temperature(self, distribution):
        """ai.py:57-66, kept for callers that use it on their own distributions."""
        log_probabilities = np.log(distribution) * self.beta
        log_probabilities = log_probabilities - log_probabilities.max()
        probabilities = np.exp(log_probabilities)
        return probabilities / probabilities.sum()

    def get_move(self, state):
        if self.move_limit is not None and _plies(state) > self.move_limit:
            return PASS_MOVE
        if self.pass_when_offered and _offered_pass(state):
            return PASS_MOVE
        greedy = self.greedy_start is not None and _plies(state) >= self.greedy_start
        return (self._greedy if greedy else self._sample).get_moves_list([state])[0]

    def get_moves(self, states):
        """Batch version of get_move. A list of moves is returned (one per state) (ai.py:107-133)"""
        if len(states) == 0:
            return []
        batch = states if isinstance(states, BatchedGameState) else BatchedGameState.from_states(states)
        plies = [0] * len(batch) if isinstance(states, BatchedGameState) else [_plies(st) for st in states]
        moves = self._sample.get_moves_list(batch)
        if self.greedy_start is not None and any(p >= self.greedy_start for p in plies):
            greedy = self._greedy.get_moves_list(batch)
            moves = [g if p >= self.greedy_start else m for m, g, p in zip(moves, greedy, plies)]
        if self.move_limit is not None:
            moves = [PASS_MOVE if p > self.move_limit else m for m, p in zip(moves, plies)]
        return moves


class MCTSPlayer(object):
    """ai.py:136-149"""

    def __init__(self, value_function, policy_function, rollout_function, lmbda=.5, c_puct=5,
                 rollout_limit=500, playout_depth=40, n_playout=100):
        self.mcts = MCTS(value_function, policy_function, rollout_function, lmbda, c_puct,
                         rollout_limit, playout_depth, n_playout)

    def get_move(self, state):
        move = self.mcts.get_move(state)         # None when no legal move is left
        if move is None:
            return PASS_MOVE
        self.mcts.update_with_move(move, size=state.size)
        return move
