"""Edge cases through the C ABI: empty batches, bad sizes, short workspaces, oversized batches, and the
empty / ragged inputs of the Python surface (policy.py:36-37,19-20,47,63; preprocessing.py:285)."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import go_oracle
from rocalphago_b200 import BatchedGameState, Preprocess, PositionView, _cabi, scoring
from rocalphago_b200.models import CNNPolicy, CNNValue

pytestmark = pytest.mark.gpu
RAG_EINVAL, RAG_ENOMEM, RAG_EUNSUPPORTED = -1, -3, -4


def _err():
    return _cabi.lib().rag_last_error().decode()


def test_empty_batches_are_accepted_everywhere_and_touch_nothing():
    L = _cabi.lib()
    src = np.arange(48, dtype=np.int32)
    sp = src.ctypes.data_as(ctypes.c_void_p)
    guard = torch.full((64,), 7, dtype=torch.int32, device="cuda")
    g = guard.data_ptr()
    assert L.rag_features_analyze(None, 0, 19, 1, None, 0, None) == 0
    assert L.rag_features_ladders(None, 0, 19, None, 0, None) == 0
    assert L.rag_features_emit(None, 0, 19, sp, 48, 0, 0, 0, None, None) == 0
    assert L.rag_features(None, 0, 19, sp, 48, 0, 0, 0, None, None, 0, None) == 0
    assert L.rag_features_static(None, 0, 19, sp, 48, 0, 0, 0, None, None) == 0
    assert L.rag_do_move(None, 0, 19, None, None, None) == 0
    assert L.rag_random_playouts(None, 0, 19, 1, 100, 8, None, 0, None) == 0
    assert L.rag_rollout_winners(None, 0, 19, 10, 1, 7.5, None, None) == 0
    assert L.rag_select_moves(None, None, 0, 19, 1.0, 1, 1, None, None) == 0
    sk = _cabi.RagSuperko()
    assert L.rag_features_analyze_superko(None, 0, 19, 1, ctypes.byref(sk), None, 0, None) == 0
    assert L.rag_features_workspace_bytes(0) >= 0 and L.rag_features_workspace_bytes(-5) == 0
    net = CNNPolicy(go_oracle.DEFAULT_FEATURES, layers=2, filters_per_layer=32, seed=1)
    h = net.model._ensure_handle(torch.device("cuda", 0))
    assert L.rag_net_forward(h, None, 0, None, None, 0, None) == 0
    assert L.rag_eval_batch(h, None, 0, sp, 48, None, None, 0, None) == 0
    assert L.rag_net_workspace_bytes(h, 0) == 0 and L.rag_eval_workspace_bytes(h, 0) == 0
    torch.cuda.synchronize()
    assert (guard == 7).all() and g == guard.data_ptr()
    # Python surface
    assert net.batch_eval_state([]) == []                                   # policy.py:36-37
    assert CNNValue(go_oracle.DEFAULT_FEATURES + ["color"], layers=2, filters_per_layer=32).batch_eval_state([]) == []
    assert Preprocess().states_to_tensor([]).shape[0] == 0
    empty = BatchedGameState(torch.zeros((0, 192), dtype=torch.int32, device="cuda"), 19)
    assert len(empty) == 0 and empty.do_move([]).numel() == 0
    assert scoring.winners(empty).shape == (0,)


def test_bad_arguments_are_refused_with_a_message():
    L = _cabi.lib()
    n = 8
    st = BatchedGameState.random_playouts(n, seed=2)
    ws = torch.empty(int(L.rag_features_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
    out = torch.empty((n, 48, 19, 19), dtype=torch.uint8, device="cuda")
    src = np.arange(48, dtype=np.int32)
    sp = src.ctypes.data_as(ctypes.c_void_p)
    for size in (0, 1, 20, 32):
        assert L.rag_features_analyze(st.packed.data_ptr(), n, size, 1, ws.data_ptr(), ws.numel(), None) in (RAG_EINVAL, RAG_EUNSUPPORTED)
        assert "size" in _err().lower() or "board" in _err().lower()
    assert L.rag_features_analyze(st.packed.data_ptr(), n, 19, 1, ws.data_ptr(), ws.numel() - 1, None) == RAG_ENOMEM
    assert "workspace" in _err()
    assert L.rag_features_analyze(None, n, 19, 1, ws.data_ptr(), ws.numel(), None) == RAG_EINVAL
    assert L.rag_features_analyze(st.packed.data_ptr(), -1, 19, 1, ws.data_ptr(), ws.numel(), None) == RAG_EINVAL
    # more positions than one launch addresses (queue items carry 21 position bits): refused before anything is read
    assert L.rag_features_analyze(st.packed.data_ptr(), (1 << 21) + 1, 19, 1, ws.data_ptr(), ws.numel(), None) == RAG_EUNSUPPORTED
    assert L.rag_features_static(st.packed.data_ptr(), L.rag_features_static_max_positions() + 1, 19, sp, 48, 0, 0, 0,
                                 out.data_ptr(), None) != 0
    bad = np.array([0, 1, 77], np.int32)
    assert L.rag_features_emit(ws.data_ptr(), n, 19, bad.ctypes.data_as(ctypes.c_void_p), 3, 0, 0, 0, out.data_ptr(), None) == RAG_EINVAL
    assert L.rag_features_emit(ws.data_ptr(), n, 19, sp, 48, 9, 0, 0, out.data_ptr(), None) == RAG_EINVAL      # unknown layout
    assert L.rag_features_emit(ws.data_ptr(), n, 19, sp, 48, 2, 1, 32, out.data_ptr(), None) == RAG_EINVAL     # conv rows: 64 or 128 channels
    assert L.rag_select_moves(out.data_ptr(), ws.data_ptr(), n, 19, 0.0, 1, 1, out.data_ptr(), None) == RAG_EINVAL
    assert "temperature" in _err()
    assert L.rag_rollout_winners(st.packed.data_ptr(), n, 19, -1, 1, 7.5, out.data_ptr(), None) == RAG_EINVAL
    with pytest.raises(ValueError):
        Preprocess(["board", "no_such_feature"])                              # preprocessing.py:285
    net = CNNPolicy(go_oracle.DEFAULT_FEATURES, layers=2, filters_per_layer=32, seed=1)
    views = st.views()
    small = PositionView(np.zeros((9, 9), np.int8), np.full((9, 9), -1, np.int16), None, 1)
    with pytest.raises(ValueError):
        net.batch_eval_state([views[0], small])                              # policy.py:39-40


def test_ragged_move_lists_and_empty_move_lists():
    net = CNNPolicy(go_oracle.DEFAULT_FEATURES, layers=2, filters_per_layer=32, seed=1)
    views = BatchedGameState.random_playouts(3, seed=4).views()
    full = net.batch_eval_state(views)
    lists = [[full[0][0][0]], [m for m, _ in full[1][:7]], []]
    got = net.batch_eval_state(views, lists)
    assert [len(r) for r in got] == [1, 7, 0]                                 # an empty list gives [] (policy.py:19-20)
    assert abs(got[0][0][1] - 1.0) < 1e-6 and abs(sum(p for _, p in got[1]) - 1.0) < 1e-5
    assert [m for m, _ in got[1]] == lists[1]
    assert net.eval_state(views[2], []) == full[2]                            # `moves or legal` (policy.py:63)
    assert net.batch_eval_state(views, []) == full                            # empty moves_lists means all legal (policy.py:47)


def test_one_position_and_odd_batch_sizes():
    """Batches that do not fill a tile, a CTA or a warp quota."""
    feats = go_oracle.DEFAULT_FEATURES
    net = CNNPolicy(feats, layers=3, filters_per_layer=64, seed=9)
    b, a, ko, pl = go_oracle.random_playouts(131, size=19, seed=21)
    want = go_oracle.features(b, a, ko, pl, feats)
    for n in (1, 2, 3, 5, 127, 131):
        batch = BatchedGameState.from_arrays(b[:n], a[:n], ko[:n], pl[:n])
        assert np.array_equal(Preprocess(feats).to_device(batch).cpu().numpy(), want[:n]), n
        probs = net.model.eval_batch_device(batch, net.preprocessor.plane_src).cpu().numpy()
        assert probs.shape == (n, 361) and np.allclose(probs.sum(1), 1.0, atol=1e-5)
        if n > 1:
            one = net.model.eval_batch_device(BatchedGameState.from_arrays(b[n - 1:n], a[n - 1:n], ko[n - 1:n], pl[n - 1:n]),
                                              net.preprocessor.plane_src).cpu().numpy()
            assert np.array_equal(one[0], probs[n - 1])                       # a position's result does not depend on its batch
