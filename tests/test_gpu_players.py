"""Batched players on the device: move selection against the reference's rules
(ai.py:33-45, 57-66, 107-133) and a device-resident self-play loop replayed on the oracle."""
import numpy as np
import pytest
import torch

from oracle import go_oracle
from rocalphago_b200 import BatchedGameState
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.players import BatchedPolicyPlayer, play_games

pytestmark = pytest.mark.gpu


def _setup(n=300, seed=5):
    b, a, ko, pl = go_oracle.random_playouts(n, size=19, seed=seed, max_ply=420, stride=4)
    net = CNNPolicy(go_oracle.DEFAULT_FEATURES, layers=3, filters_per_layer=32, dtype="fp16", seed=9)
    w = net.model.get_weights()
    w[-3] = w[-3] * 300.0                                # a peaked distribution makes the checks sharper
    net.model.set_weights(w)
    batch = BatchedGameState.from_arrays(b, a, ko, pl)
    probs = net.model.eval_batch_device(batch, net.preprocessor.plane_src).cpu().numpy()
    lists, n_noneye = go_oracle.legal_moves(b, ko, pl)
    sensible = [m[:k] for m, k in zip(lists, n_noneye)]
    return net, batch, probs, sensible


def test_greedy_player_picks_the_most_probable_sensible_move():
    net, batch, probs, sensible = _setup()
    mv = BatchedPolicyPlayer(net, greedy=True).get_moves(batch).cpu().numpy()
    for i, s in enumerate(sensible):
        if len(s) == 0:
            assert mv[i] == -1
        else:
            assert mv[i] in set(s.tolist())
            assert probs[i][mv[i]] == probs[i][s].max()
            assert mv[i] == s[np.argmax(probs[i][s])]     # first maximum in move-list (x-major) order
    tup = BatchedPolicyPlayer(net, greedy=True).get_moves_list(batch)
    assert tup[0] is None or tup[0] == (int(mv[0]) // 19, int(mv[0]) % 19)


@pytest.mark.parametrize("temperature", [1.0, 0.5, 2.0])
def test_probabilistic_player_draws_from_the_tempered_distribution(temperature):
    """apply_temperature (ai.py:57-66): p^(1/T) over the sensible moves, renormalised.  4096 copies of
    one position give 4096 independent draws; their frequencies must match that distribution."""
    net, batch, probs, sensible = _setup(n=8, seed=11)
    k = 3
    copies = BatchedGameState(batch.packed[k:k + 1].repeat(4096, 1).contiguous(), 19)
    player = BatchedPolicyPlayer(net, temperature=temperature, seed=123)
    draws = torch.cat([player.get_moves(copies) for _ in range(4)]).cpu().numpy()
    s = sensible[k]
    assert set(draws.tolist()) <= set(s.tolist())
    p = probs[k][s].astype(np.float64) ** (1.0 / temperature)
    p /= p.sum()
    freq = np.asarray([(draws == m).mean() for m in s])
    # 16384 draws: standard error of a frequency <= 0.004
    assert np.abs(freq - p).max() < 0.02, float(np.abs(freq - p).max())
    assert np.abs(freq - p).sum() < 0.25
    # greedy_start switches to the arg-max
    g = BatchedPolicyPlayer(net, temperature=temperature, greedy_start=10).get_moves(copies, ply=10).cpu().numpy()
    assert (g == s[np.argmax(probs[k][s])]).all()
    lim = BatchedPolicyPlayer(net, move_limit=5).get_moves(copies, ply=6).cpu().numpy()
    assert (lim == -1).all()


def test_device_self_play_replays_on_the_oracle():
    """64 games of 120 plies played entirely on the GPU; the oracle replays every move list and
    reaches the same final boards (every selected move was legal and sensible at its time)."""
    net = CNNPolicy(go_oracle.DEFAULT_FEATURES, layers=2, filters_per_layer=32, dtype="fp16", seed=2)
    final, log = play_games(BatchedPolicyPlayer(net, temperature=1.0, seed=77), 64, max_moves=120, record=True)
    log = log.cpu().numpy()
    gb, ga, gko, gpl = final.to_arrays()
    for gme in range(64):
        played, rb, ra, rko, rpl = go_oracle.replay(19, log[:, gme], None, every=False)
        assert played == log.shape[0]
        assert np.array_equal(rb[0], gb[gme]) and rko[0] == gko[gme] and rpl[0] == gpl[gme]
        assert np.array_equal(np.minimum(ra[0], 7), ga[gme])
    assert (gb != 0).sum() > 64 * 80


def test_reference_named_players_and_search():
    """ai.py's class names and constructor arguments: GreedyPolicyPlayer / ProbabilisticPolicyPlayer /
    MCTSPlayer, and mcts.MCTS(value_fn, policy_fn, rollout_policy_fn, ...) taking bound eval_state methods."""
    from rocalphago_b200 import PositionView
    from rocalphago_b200.ai import GreedyPolicyPlayer, MCTSPlayer, PASS_MOVE, ProbabilisticPolicyPlayer
    from rocalphago_b200.mcts import MCTS
    from rocalphago_b200.models import CNNValue
    net, batch, probs, sensible = _setup(n=40)
    b, a, ko, pl = batch.to_arrays()

    class State(PositionView):                                   # a GameState look-alike with a move history
        def __init__(self, i, plies):
            PositionView.__init__(self, b[i], a[i], int(ko[i]), int(pl[i]))
            self.history = [(0, 0)] * plies

    states = [State(i, plies=10 * i) for i in range(40)]
    greedy = GreedyPolicyPlayer(net, move_limit=300)
    for i in (0, 7, 31, 39):
        mv = greedy.get_move(states[i])
        if 10 * i > 300 or len(sensible[i]) == 0:
            assert mv is PASS_MOVE
        else:
            s = sensible[i]
            assert mv == (int(s[np.argmax(probs[i][s])]) // 19, int(s[np.argmax(probs[i][s])]) % 19)
    prob = ProbabilisticPolicyPlayer(net, temperature=0.7, move_limit=350, greedy_start=200, seed=3)
    moves = prob.get_moves(states)
    assert len(moves) == 40
    for i, mv in enumerate(moves):
        s = sensible[i]
        if 10 * i > 350 or len(s) == 0:
            assert mv is PASS_MOVE
        elif 10 * i >= 200:
            assert mv == (int(s[np.argmax(probs[i][s])]) // 19, int(s[np.argmax(probs[i][s])]) % 19)
        else:
            assert mv[0] * 19 + mv[1] in set(s.tolist())
    assert prob.get_move(states[25]) == moves[25]                # greedy past greedy_start
    assert abs(prob.apply_temperature(np.array([0.2, 0.3, 0.5])).sum() - 1.0) < 1e-12
    value = CNNValue(go_oracle.DEFAULT_FEATURES + ["color"], layers=2, filters_per_layer=32, dense_units=16, seed=4)
    search = MCTS(value.eval_state, net.eval_state, net.eval_state, lmbda=0.0, c_puct=5, playout_depth=4, n_playout=12)
    mv = search.get_move(states[3])
    assert mv is None or mv[0] * 19 + mv[1] in set(go_oracle.legal_moves(b[3:4], ko[3:4], pl[3:4])[0][0].tolist())
    player = MCTSPlayer(value.eval_state, net.eval_state, net.eval_state, lmbda=0.0, playout_depth=3, n_playout=8)
    assert player.get_move(states[5]) is not None
    with pytest.raises(TypeError):
        MCTS(lambda st: 0.0, net.eval_state, None)
