"""Positional superko (go.py:242-265, 282-283) on the device, through the C ABI: planes, legal-move lists
and network evaluations of positions recorded from reference games played with
GameState(enforce_superko=True) (tests/golden/features_superko.npz, oracle/make_golden_superko.py)."""
import numpy as np
import pytest
import torch

from oracle import go_oracle
from rocalphago_b200 import BatchedGameState, Preprocess
from rocalphago_b200.models import CNNPolicy
from rocalphago_b200.state import SuperkoData
from tests import golden_util as gu

pytestmark = pytest.mark.gpu


class _State(object):
    """What a reference GameState created with enforce_superko=True exposes to this path."""

    def __init__(self, g, i, size):
        P = size * size
        self.size = size
        self.board, self.stone_ages = g["board"][i], g["age"][i]
        self.ko = None if g["ko"][i] < 0 else (int(g["ko"][i]) // size, int(g["ko"][i]) % size)
        self.current_player = int(g["player"][i])
        self.enforce_superko = True
        m = g["masks"][i].reshape(3, P)
        even = [(p // size, p % size) for p in np.flatnonzero(m[0])]
        odd = [(p // size, p % size) for p in np.flatnonzero(m[1])]
        n = int(g["hist_len"][i])
        hist = [None] * n                                   # passes fill what the masks do not account for
        hist[0:2 * len(even):2] = even
        hist[1:2 * len(odd):2] = odd
        assert len(hist) == n
        self.history = hist
        self.handicaps = [(p // size, p % size) for p in np.flatnonzero(m[2])]
        self.current_hash = np.uint64(g["hash"][i])
        self.previous_hashes = set(int(h) for h in g["prev"][g["prev_off"][i]:g["prev_off"][i + 1]])


def _batch(g, size):
    b = BatchedGameState.from_arrays(g["board"], g["age"], g["ko"], g["player"])
    sk = g["sk"]
    b.superko = SuperkoData.from_arrays(size, b.device, sk["enforce"], sk["has_handicap"], sk["hist_len"], sk["hash"],
                                        sk["prev"], sk["prev_off"], sk["masks"])
    return b


def test_planes_and_move_lists_with_superko_match_the_reference():
    feats, sets = gu.load_superko()
    pp = Preprocess(feats)
    changed = 0
    for key, g in sets.items():
        size = g["board"].shape[-1]
        batch = _batch(g, size)
        got = pp.to_device(batch).cpu().numpy()
        assert np.array_equal(got, g["planes"]), key
        lists = pp.legal_move_lists(pp.analyze(batch), len(batch), size)
        for i, mv in enumerate(lists):
            want = g["moves"][i][g["moves"][i] >= 0]
            assert [x * size + y for x, y in mv] == want.tolist(), (key, i)
        plain = pp.to_device(BatchedGameState.from_arrays(g["board"], g["age"], g["ko"], g["player"])).cpu().numpy()
        changed += int((plain != got).reshape(len(got), -1).any(1).sum())
        # positions that do not enforce it are evaluated as before, inside the same call
        off = _batch(dict(g, sk=dict(g["sk"], enforce=np.zeros(len(got), np.int32))), size)
        assert np.array_equal(pp.to_device(off).cpu().numpy(), plain), key
    assert changed >= 150


def test_gamestate_objects_with_enforce_superko_are_honoured_end_to_end():
    """from_states reads enforce_superko / history / handicaps / current_hash / previous_hashes;
    batch_eval_state then lists only the moves the reference's get_legal_moves() gives."""
    feats, sets = gu.load_superko()
    net = CNNPolicy(feats[:-1], board=3, layers=2, filters_per_layer=32, seed=5)          # the 48 default planes
    g = sets["3_0"]
    pick = [i for i in range(len(g["board"])) if (g["planes"][i, 48] != go_oracle.features(
        g["board"][i:i + 1], g["age"][i:i + 1], g["ko"][i:i + 1], g["player"][i:i + 1], ["legal"])[0, 0]).any()][:24]
    assert len(pick) >= 10                                   # positions where superko forbids something
    states = [_State(g, i, 3) for i in pick]
    out = net.batch_eval_state(states)
    for i, res in zip(pick, out):
        want = g["moves"][i][g["moves"][i] >= 0].tolist()
        assert [x * 3 + y for (x, y), _ in res] == want
        if want:
            assert abs(sum(p for _, p in res) - 1.0) < 1e-5
    # the same states without the flag list more moves
    for st in states:
        st.enforce_superko = False
    more = net.batch_eval_state(states)
    assert sum(len(r) for r in more) > sum(len(r) for r in out)
    # Preprocess.state_to_tensor on one such state
    states[0].enforce_superko = True
    t = Preprocess(feats).state_to_tensor(states[0])
    assert np.array_equal(t[0].astype(np.uint8), g["planes"][pick[0]])
    # a device do_move drops the history: the batch no longer enforces superko
    b = BatchedGameState.from_states(states[:1])
    assert b.superko is not None
    b.do_move([-1])
    assert b.superko is None


def test_superko_entry_point_rejects_bad_arguments():
    import ctypes
    from rocalphago_b200 import _cabi
    L = _cabi.lib()
    ws = torch.empty(int(L.rag_features_workspace_bytes(4)), dtype=torch.uint8, device="cuda")
    st = torch.zeros((4, 192), dtype=torch.int32, device="cuda")
    assert L.rag_features_analyze_superko(st.data_ptr(), 4, 19, 1, None, ws.data_ptr(), ws.numel(), None) != 0
    assert b"superko" in L.rag_last_error()
    empty = _cabi.RagSuperko()
    assert L.rag_features_analyze_superko(st.data_ptr(), 4, 19, 1, ctypes.byref(empty), ws.data_ptr(), ws.numel(), None) != 0
    assert L.rag_features_analyze_superko(st.data_ptr(), 4, 19, 2, ctypes.byref(empty), ws.data_ptr(), ws.numel(), None) != 0
