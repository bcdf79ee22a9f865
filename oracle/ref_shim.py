"""Imports the UNMODIFIED reference (hiropppe/RocAlphaGo) from /root/reference -- TEST INFRASTRUCTURE ONLY.

Works only where /root/reference is mounted (the build container), never on the
GPU box.  Used by oracle/make_golden.py to produce tests/golden/* and by
oracle/check_against_reference.py to validate go_oracle.c at scale.

The reference targets Python 2.7 / numpy 1.11 / Keras 1.2.0; two shims let its
rules engine and feature code import under Python 3.12:
`np.int`/`np.float` aliases (go.py:44,55) and a stub `keras.backend`
(preprocessing.py:3,7).  No reference file is copied or edited.
"""
import os
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("RAG_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "AlphaGo"))


def load():
    """Returns (go_module, preprocessing_module) of the reference."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    if not hasattr(np, "int"):
        np.int = int
    if not hasattr(np, "float"):
        np.float = float
    if "keras" not in sys.modules:
        keras = types.ModuleType("keras")
        kb = types.ModuleType("keras.backend")
        kb.set_image_dim_ordering = lambda _order: None
        keras.backend = kb
        sys.modules["keras"] = keras
        sys.modules["keras.backend"] = kb
    for p in (os.path.join(REFERENCE_ROOT, "tests"), REFERENCE_ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    from AlphaGo import go
    from AlphaGo.preprocessing import preprocessing
    return go, preprocessing


def state_arrays(state):
    """(board int8 (S,S), age int16 (S,S), ko int, player int) of a reference GameState."""
    S = state.size
    ko = -1 if state.ko is None else state.ko[0] * S + state.ko[1]
    return (np.asarray(state.board, dtype=np.int8).copy(),
            np.asarray(state.stone_ages, dtype=np.int16).copy(), int(ko), int(state.current_player))


def state_from_arrays(go, board, age, ko, player):
    """Rebuilds a reference GameState whose derived sets match `board`.

    Stones are placed with do_move in any order -- a legal position never captures
    while it is rebuilt -- then ages / ko / side to move are overwritten.
    """
    S = board.shape[-1]
    st = go.GameState(size=S)
    for x in range(S):
        for y in range(S):
            if board[x, y] != 0:
                st.do_move((x, y), int(board[x, y]))
    assert np.array_equal(st.board, board), "position is not capture-free to rebuild"
    st.stone_ages = np.asarray(age, dtype=int).copy()
    st.ko = None if ko < 0 else (int(ko) // S, int(ko) % S)
    st.current_player = int(player)
    st._GameState__legal_move_cache = None
    st._GameState__legal_eyes_cache = None
    st.history = []
    return st
