"""Host-side logic of the reference-facing layer that needs no GPU: packing of GameState fields (single- and
multi-threaded), the lazy move lists batch_eval_state hands back, SGF points off the board, dtype names."""
import ctypes

import numpy as np
import pytest

from oracle import go_oracle
from rocalphago_b200 import PositionView, _cabi
from rocalphago_b200.game_converter import _flat
from rocalphago_b200.models.policy import MoveDistributions, MoveList
from rocalphago_b200.preprocessing import DTYPE_CODES
from rocalphago_b200.state import arrays_from_states, pack_arrays


def _numpy_pack(board, age, ko, player):
    """The packed-state layout of include/rocalphago_b200.h written out with numpy."""
    n, S = board.shape[0], board.shape[-1]
    out = np.zeros((n, 192), np.uint32)
    bit = (1 << np.arange(S, dtype=np.uint32))[None, None, :]
    a = np.clip(age, 0, 7)
    stone = board != 0
    out[:, 0:S] = ((board == 1) * bit).sum(-1)
    out[:, 32:32 + S] = ((board == -1) * bit).sum(-1)
    for k, base in enumerate((64, 96, 128)):
        out[:, base:base + S] = ((stone & (((a >> k) & 1) == 1)) * bit).sum(-1)
    out[:, 168:168 + S] = ((stone & (age < 0)) * bit).sum(-1)
    out[:, 160] = np.where(ko < 0, -1, ko).astype(np.int64).astype(np.uint32)
    out[:, 161] = player.astype(np.int64).astype(np.uint32)
    out[:, 162] = S
    return out


@pytest.mark.parametrize("n,size", [(3, 19), (1500, 19), (1100, 9)])
def test_packing_matches_the_layout_single_and_multi_threaded(n, size):
    b, a, ko, pl = go_oracle.random_playouts(n, size=size, seed=n)
    a = a.copy()
    a[::7][b[::7] != 0] = -1                                       # stones GameState.copy() left without an age
    packed = pack_arrays(b, a, ko, pl).numpy().view(np.uint32)
    assert np.array_equal(packed, _numpy_pack(b.astype(np.int64), a.astype(np.int64), ko, pl))
    # and back
    L = _cabi.lib()
    B, A, K, P = np.zeros_like(b), np.zeros_like(a), np.zeros(n, np.int32), np.zeros(n, np.int32)
    vp = ctypes.c_void_p
    assert L.rag_unpack_states(n, size, vp(packed.ctypes.data), vp(B.ctypes.data), vp(A.ctypes.data), vp(K.ctypes.data), vp(P.ctypes.data)) == 0
    assert np.array_equal(B, b) and np.array_equal(K, ko) and np.array_equal(P, pl)
    assert np.array_equal(A[b != 0], np.where(a < 0, -1, np.minimum(a, 7))[b != 0])


def test_bad_fields_are_refused_whatever_thread_meets_them():
    b, a, ko, pl = go_oracle.random_playouts(2000, size=19, seed=5)
    for where in (0, 777, 1999):
        bad = b.copy()
        bad[where, 3, 3] = 2
        with pytest.raises(_cabi.RagError, match=r"board\[%d\]" % where):
            pack_arrays(bad, a, ko, pl)
    k2 = ko.copy()
    k2[1234] = 361
    with pytest.raises(_cabi.RagError, match=r"ko\[1234\]"):
        pack_arrays(b, a, k2, pl)


def test_states_are_gathered_like_np_stack_would():
    b, a, ko, pl = go_oracle.random_playouts(40, size=13, seed=9)
    states = [PositionView(b[i].astype(np.float64), a[i].astype(np.int64), int(ko[i]), int(pl[i])) for i in range(40)]
    B, A, K, P = arrays_from_states(states)
    assert B.dtype == np.int8 and A.dtype == np.int16
    assert np.array_equal(B, b) and np.array_equal(A, a) and np.array_equal(K, ko) and np.array_equal(P, pl)
    with pytest.raises(ValueError, match="same size"):
        arrays_from_states(states + [PositionView(np.zeros((9, 9)), -np.ones((9, 9)), None, 1)])


def test_move_lists_behave_like_the_references_lists():
    S = 5
    moves = np.array([[7, 0, 24, 3] + [0] * 21, [1] + [0] * 24, [0] * 25], np.int16)
    probs = np.zeros((3, 25), np.float32)
    probs[0, :4] = [0.5, 0.25, 0.125, 0.125]
    probs[1, 0] = 1.0
    res = MoveDistributions(moves, probs, np.array([4, 1, 0], np.int32), S).as_list()
    assert isinstance(res, list) and len(res) == 3 and all(isinstance(r, MoveList) for r in res)
    assert [len(r) for r in res] == [4, 1, 0]
    assert res[0] == [((1, 2), 0.5), ((0, 0), 0.25), ((4, 4), 0.125), ((0, 3), 0.125)] and res[2] == [] and res[1] != []
    assert res[0][1] == ((0, 0), 0.25) and res[0][-1] == ((0, 3), 0.125) and res[0][1:3] == [((0, 0), 0.25), ((4, 4), 0.125)]
    m, p = zip(*res[0])                                            # ai.py:121
    assert m == ((1, 2), (0, 0), (4, 4), (0, 3)) and sum(p) == 1.0
    assert max(res[0], key=lambda mp: mp[1])[0] == (1, 2)          # ai.py:35
    assert [mv for mv, _ in res[1]] == [(0, 1)] and list(res[2]) == [] and res[0] == res[0].to_list() == list(res[0])
    assert repr(res[1]) == "[((0, 1), 1.0)]"
    with pytest.raises(IndexError):
        res[1][1]


def test_sgf_points_past_the_board_do_not_alias_other_points():
    assert _flat((0, 0), 9) == 0 and _flat((8, 8), 9) == 80 and _flat(None, 9) == -1
    assert _flat((5, 25), 19) == 19 * 19 and _flat((25, 0), 19) == 19 * 19 and _flat((0, 9), 9) == 81


def test_dtype_names():
    assert DTYPE_CODES == {"bf16": 0, "fp16": 1, "fp16x3": 2, "fp16c32": 3, "fp16c64": 4}
    from rocalphago_b200.models.nn_util import Network
    with pytest.raises(ValueError, match="dtype"):
        Network(19, 48, [(32, 5)], dtype="fp32")
    from rocalphago_b200.models.policy import CNNPolicy, DEFAULT_POLICY_DTYPE
    assert CNNPolicy.create_network(input_dim=48, layers=2, filters_per_layer=32).dtype == DEFAULT_POLICY_DTYPE == "fp16c32"
