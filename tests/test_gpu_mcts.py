"""Batched tree search (rocalphago_b200/mcts.py) and device rollouts / scoring, SURVEY 8f-3.

* with batch=1 the search must be the reference's MCTS playout for playout: visit counts, Q values and
  chosen moves recorded from the reference's own mcts.py (oracle/make_golden_mcts.py) on its GameState,
  with CPU-oracle networks, against the device networks (dtype fp16x3) -- also after update_with_move;
* get_winner and random rollouts against the reference's get_winner() values and the C oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import go_oracle
from rocalphago_b200 import BatchedGameState, scoring
from rocalphago_b200.mcts import BatchedMCTS
from rocalphago_b200.models import CNNPolicy, CNNValue
from tests import golden_util as gu

pytestmark = pytest.mark.gpu
FEATS = go_oracle.DEFAULT_FEATURES


def _nets(z, dtype="fp16x3"):
    S, layers, filters = int(z["size"]), int(z["layers"]), int(z["filters"])
    pol = CNNPolicy(FEATS, board=S, layers=layers, filters_per_layer=filters, dtype=dtype)
    pol.model.set_weights([z["pw%02d" % i] for i in range(2 * layers + 3)])
    val = CNNValue(FEATS + ["color"], board=S, layers=layers, filters_per_layer=filters, dense_units=64, dtype=dtype)
    val.model.set_weights([z["vw%02d" % i] for i in range(2 * layers + 6)])
    return pol, val


def _position(size, moves):
    _, b, a, ko, pl = go_oracle.replay(size, moves)
    return BatchedGameState.from_arrays(b, a, ko, pl)


@pytest.mark.parametrize("case", ["open", "fight"])
def test_batch_one_search_is_the_references_search(case):
    z = np.load(os.path.join(gu.GOLDEN, "mcts_reference.npz"))
    S = int(z["size"])
    pol, val = _nets(z)
    n_playout, depth, c_puct = [int(v) for v in z["%s_params" % case]]
    setup = z["%s_setup" % case].tolist()
    search = BatchedMCTS(pol, val, lmbda=0, c_puct=c_puct, playout_depth=depth, n_playout=n_playout, batch=1)
    move = search.get_move(_position(S, setup))
    root = search._root
    assert root.actions.tolist() == z["%s_actions" % case].tolist()
    assert root.n.tolist() == z["%s_visits" % case].tolist()
    assert np.allclose(root.Q, z["%s_q" % case], atol=2e-5)
    assert move[0] * S + move[1] == int(z["%s_move" % case]) and root.visits == n_playout
    # re-root on the chosen move and search again from the new position, as a game loop does
    search.update_with_move(move, size=S)
    move2 = search.get_move(_position(S, setup + [move[0] * S + move[1]]))
    assert search._root.actions.tolist() == z["%s_actions2" % case].tolist()
    assert search._root.n.tolist() == z["%s_visits2" % case].tolist()
    assert move2[0] * S + move2[1] == int(z["%s_move2" % case])


def test_batched_waves_keep_the_tree_consistent():
    z = np.load(os.path.join(gu.GOLDEN, "mcts_reference.npz"))
    S = int(z["size"])
    pol, val = _nets(z, dtype="fp16")
    pos = _position(S, z["fight_setup"].tolist())
    results = []
    for batch in (8, 32):
        search = BatchedMCTS(pol, val, c_puct=2, playout_depth=6, n_playout=160, batch=batch)
        move = search.get_move(pos)
        root = search._root
        assert root.visits == 160 and int(root.n.sum()) == 160            # every playout passes through one child of the root
        assert (root.virtual == 0).all()                                   # virtual visits are all taken back
        assert move == max(search.visit_counts(S).items(), key=lambda kv: kv[1])[0]

        def check(node):
            if node.is_leaf():
                return
            assert (node.virtual == 0).all() and (node.n >= 0).all()
            for i, ch in enumerate(node.children):
                if ch is not None and not ch.is_leaf():
                    assert int(ch.n.sum()) <= int(node.n[i])               # visits of a child's children never exceed the child's
                    check(ch)
        check(root)
        results.append(search.visit_counts(S))
        assert search.evaluations < 160 * 6                                 # shared leaves are expanded once per wave
    # same inputs, same schedule -> same tree
    again = BatchedMCTS(pol, val, c_puct=2, playout_depth=6, n_playout=160, batch=32)
    again.get_move(pos)
    assert again.visit_counts(S) == results[1]


def test_rollouts_in_the_leaf_value():
    z = np.load(os.path.join(gu.GOLDEN, "mcts_reference.npz"))
    S = int(z["size"])
    pol, val = _nets(z, dtype="fp16")
    pos = _position(S, z["open_setup"].tolist())
    search = BatchedMCTS(pol, val, lmbda=0.5, c_puct=5, playout_depth=3, n_playout=64, batch=16, rollout_limit=200, seed=3)
    move = search.get_move(pos)
    assert move is not None and search._root.visits == 64
    assert -1.0 <= search._root.q <= 1.0
    only = BatchedMCTS(pol, None, lmbda=1.0, playout_depth=2, n_playout=32, batch=16, rollout_limit=200)
    only.get_move(pos)
    assert set(np.unique(np.round(only._root.Q[only._root.n > 0] * only._root.n[only._root.n > 0], 6) % 1)) <= {0.0}   # sums of +-1 / 0
    with pytest.raises(ValueError):
        BatchedMCTS(pol, None, lmbda=0.5)


def test_winners_match_the_reference_and_rollouts_match_the_oracle():
    w = np.load(os.path.join(gu.GOLDEN, "winners.npz"))
    for size in (19, 9, 5):
        g = gu.load_features("features_playouts_%d.npz" % size)
        r = w["rows_%d" % size]
        batch = BatchedGameState.from_arrays(g["board"][r], g["age"][r], g["ko"][r], g["player"][r])
        for komi in (7.5, 0.0):
            assert np.array_equal(scoring.winners(batch, komi), w["winner_%d_%g" % (size, komi)]), (size, komi)
    for size, n in ((9, 400), (19, 300), (5, 200)):
        b, a, ko, pl = go_oracle.random_playouts(n, size=size, seed=77 + size, max_ply=size * size, stride=3)
        batch = BatchedGameState.from_arrays(b, a, ko, pl)
        before = batch.packed.clone()
        for limit, seed in ((500, 5), (7, 6)):
            got = scoring.random_rollout_results(batch, limit=limit, seed=seed)
            want = go_oracle.rollout_winners(b, a, ko, pl, limit=limit, seed=seed)
            assert np.array_equal(got.astype(np.int32), want), (size, limit)
        assert torch.equal(batch.packed, before)                              # inputs are not modified


def test_parallel_search_builds_the_same_trees_as_searching_each_position_alone():
    """ParallelMCTS advances one tree per position in lock step (one policy batch per ply over all trees); since
    the device evaluates every position independently of its batch, each tree must equal the single search's."""
    from rocalphago_b200.mcts import ParallelMCTS
    z = np.load(os.path.join(gu.GOLDEN, "mcts_reference.npz"))
    S = int(z["size"])
    pol, val = _nets(z, dtype="fp16")
    setups = [z["fight_setup"].tolist(), z["open_setup"].tolist(), z["fight_setup"].tolist()[:-2]]
    roots = BatchedGameState(torch.cat([_position(S, s).packed for s in setups]), S)
    kw = dict(c_puct=2, playout_depth=5, n_playout=96, batch=16)
    par = ParallelMCTS(pol, val, games=3, **kw)
    moves = par.get_moves(roots)
    for g, setup in enumerate(setups):
        alone = BatchedMCTS(pol, val, **kw)
        mv = alone.get_move(_position(S, setup))
        assert mv == moves[g]
        assert alone.visit_counts(S) == par.searches[g].visit_counts(S)
        assert np.array_equal(alone._root.Q, par.searches[g]._root.Q)
    chosen = [par.searches[g].visit_counts(S)[moves[g]] for g in range(3)]
    par.update_with_moves(moves, size=S)
    assert [s._root.visits for s in par.searches] == chosen           # every tree re-rooted on its own move
