"""CPU-only checks of the C-ABI library: it loads, exports every symbol the public header
declares, and its host-side entry points (packing) work.  No compute kernel is called."""
import ctypes

import numpy as np
import pytest

from oracle import go_oracle
from rocalphago_b200 import _cabi, state


def test_library_exports_every_declared_symbol():
    L = _cabi.lib()
    names = _cabi._declared_symbols()
    assert len(names) >= 12
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, "declared in include/rocalphago_b200.h but not exported: %s" % missing
    assert L.rag_abi_version() >= 1


def test_pack_unpack_roundtrip():
    for size in (19, 13, 9, 5):
        b, a, ko, pl = go_oracle.random_playouts(50, size=size, seed=size, max_ply=size * size, stride=3)
        packed = state.pack_arrays(b, a, ko, pl)
        assert tuple(packed.shape) == (50, 192)
        b2, a2, ko2, pl2 = state.BatchedGameState(packed, size).to_arrays()
        assert np.array_equal(b, b2) and np.array_equal(ko, ko2) and np.array_equal(pl, pl2)
        assert np.array_equal(np.minimum(a, 7), a2)


def test_pack_rejects_bad_input():
    L = _cabi.lib()
    board = np.zeros((1, 19 * 19), np.int8)
    board[0, 5] = 3
    age = np.zeros((1, 361), np.int16)
    ko = np.asarray([-1], np.int32)
    pl = np.asarray([1], np.int32)
    out = np.zeros(192, np.uint32)
    vp = lambda x: x.ctypes.data_as(ctypes.c_void_p)  # noqa: E731
    assert L.rag_pack_states(1, 19, vp(board), vp(age), vp(ko), vp(pl), vp(out)) == -1
    assert b"board" in L.rag_last_error()
    assert L.rag_pack_states(1, 25, vp(board), vp(age), vp(ko), vp(pl), vp(out)) == -4
    with pytest.raises(_cabi.RagError):
        _cabi.check(L.rag_pack_states(1, 25, vp(board), vp(age), vp(ko), vp(pl), vp(out)))


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    b, a, ko, pl = go_oracle.random_playouts(2, seed=1)
    with pytest.raises(_cabi.RagError):
        state.BatchedGameState.from_arrays(b, a, ko, pl)


def test_preprocess_registry_matches_reference_names():
    from rocalphago_b200.preprocessing import Preprocess, DEFAULT_FEATURES, FEATURES
    pp = Preprocess()
    assert pp.output_dim == 48 and pp.feature_list == DEFAULT_FEATURES
    assert Preprocess(["Board", "LEGAL"]).output_dim == 4
    assert set(FEATURES) == set(go_oracle.FEATURE_IDS)
    assert set(FEATURES) - {"color"} == {"board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
                                          "liberties_after", "ladder_capture", "ladder_escape", "sensibleness", "zeros", "legal"}
    with pytest.raises(ValueError, match="uknown feature"):
        Preprocess(["board", "nonsense"])
