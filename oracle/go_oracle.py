"""ctypes front end of oracle/go_oracle.c -- TEST INFRASTRUCTURE ONLY.

The C file is the CPU restatement of the reference's rules engine
(`AlphaGo/go.py`) and feature functions (`AlphaGo/preprocessing/preprocessing.py`);
this module only marshals numpy arrays into it.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import it; nothing under `rocalphago_b200/` does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libgo_oracle.so")

# feature ids, same order as preprocessing.py:209-258's FEATURES keys
FEATURE_IDS = {
    "board": 0, "ones": 1, "turns_since": 2, "liberties": 3, "capture_size": 4,
    "self_atari_size": 5, "liberties_after": 6, "ladder_capture": 7, "ladder_escape": 8,
    "sensibleness": 9, "zeros": 10, "legal": 11, "color": 12,
}
DEFAULT_FEATURES = [
    "board", "ones", "turns_since", "liberties", "capture_size", "self_atari_size",
    "liberties_after", "ladder_capture", "ladder_escape", "sensibleness", "zeros"]

_lib = None


def build(force=False):
    """Compile go_oracle.c with gcc (oracle/Makefile)."""
    src = os.path.join(_HERE, "go_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
        i16p = np.ctypeslib.ndpointer(np.int16, flags="C")
        i8p = np.ctypeslib.ndpointer(np.int8, flags="C")
        u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
        L.og_features.restype = ctypes.c_int
        L.og_features.argtypes = [ctypes.c_int, ctypes.c_int, i8p, i16p, i32p, i32p, i32p, ctypes.c_int, u8p]
        L.og_legal_moves_batch.restype = None
        L.og_legal_moves_batch.argtypes = [ctypes.c_int, ctypes.c_int, i8p, i32p, i32p, i16p, i32p]
        L.og_replay.restype = ctypes.c_int
        L.og_replay.argtypes = [ctypes.c_int, ctypes.c_int, i32p, ctypes.c_void_p, ctypes.c_int,
                                i8p, i16p, i32p, i32p]
        L.og_random_playouts.restype = ctypes.c_int
        L.og_random_playouts.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
                                         ctypes.c_int, i8p, i16p, i32p, i32p, ctypes.c_void_p]
        L.og_do_move_batch.restype = None
        L.og_do_move_batch.argtypes = [ctypes.c_int, ctypes.c_int, i8p, i16p, i32p, i32p, i32p, i32p]
        L.og_ladder_query.restype = ctypes.c_int
        L.og_ladder_query.argtypes = [ctypes.c_int, i8p] + [ctypes.c_int] * 5
        L.og_feature_planes.restype = ctypes.c_int
        L.og_feature_planes.argtypes = [ctypes.c_int]
        _lib = L
    return _lib


def _canon(board, age, ko, player):
    board = np.ascontiguousarray(board, dtype=np.int8)
    n = board.shape[0]
    size = board.shape[-1]
    board = board.reshape(n, size * size)
    age = np.ascontiguousarray(age, dtype=np.int16).reshape(n, size * size)
    ko = np.ascontiguousarray(ko, dtype=np.int32).reshape(n)
    player = np.ascontiguousarray(player, dtype=np.int32).reshape(n)
    return n, size, board, age, ko, player


def features(board, age, ko, player, feature_list=DEFAULT_FEATURES):
    """uint8 (n, F, S, S) planes for packed states.

    board (n,S,S) int8 in {+1,-1,0}; age (n,S,S) int16 (-1 on empties);
    ko (n,) int32 = x*S+y or -1; player (n,) int32 = +1/-1.
    """
    n, size, board, age, ko, player = _canon(board, age, ko, player)
    ids = np.asarray([FEATURE_IDS[f.lower()] for f in feature_list], dtype=np.int32)
    F = int(sum(lib().og_feature_planes(int(i)) for i in ids))
    out = np.empty((n, F, size, size), dtype=np.uint8)
    got = lib().og_features(n, size, board, age, ko, player, ids, len(ids), out.reshape(-1))
    assert got == F
    return out


def legal_moves(board, ko, player):
    """Per state: flat move indices in get_legal_moves() order, and the count of non-eye moves."""
    board = np.ascontiguousarray(board, dtype=np.int8)
    n, size = board.shape[0], board.shape[-1]
    moves = np.empty((n, size * size), dtype=np.int16)
    counts = np.empty((n, 2), dtype=np.int32)
    lib().og_legal_moves_batch(n, size, board.reshape(n, -1), np.ascontiguousarray(ko, np.int32),
                               np.ascontiguousarray(player, np.int32), moves, counts)
    return [moves[i, :counts[i, 0]].astype(np.int64) for i in range(n)], counts[:, 1].copy()


def replay(size, moves, colors=None, every=False):
    """Play a move list (flat indices, -1 = pass) from the empty board."""
    moves = np.ascontiguousarray(moves, dtype=np.int32)
    k = len(moves) if every else 1
    board = np.zeros((max(k, 1), size * size), np.int8)
    age = np.zeros((max(k, 1), size * size), np.int16)
    ko = np.zeros(max(k, 1), np.int32)
    player = np.zeros(max(k, 1), np.int32)
    cptr = None
    if colors is not None:
        colors = np.ascontiguousarray(colors, dtype=np.int32)
        cptr = colors.ctypes.data_as(ctypes.c_void_p)
    played = lib().og_replay(size, len(moves), moves, cptr, int(every), board, age, ko, player)
    shape = (-1, size, size)
    return played, board.reshape(shape), age.reshape(shape), ko, player


def random_playouts(n, size=19, seed=20261106, max_ply=400, stride=8, with_moves=False):
    """n positions sampled from seeded random playouts."""
    board = np.zeros((n, size * size), np.int8)
    age = np.zeros((n, size * size), np.int16)
    ko = np.zeros(n, np.int32)
    player = np.zeros(n, np.int32)
    moves = np.zeros(n, np.int32)
    got = lib().og_random_playouts(size, seed, max_ply, stride, n, board, age, ko, player,
                                   moves.ctypes.data_as(ctypes.c_void_p))
    assert got == n
    out = (board.reshape(n, size, size), age.reshape(n, size, size), ko, player)
    return out + (moves,) if with_moves else out


def do_move_batch(board, age, ko, player, moves):
    """Apply one move per state (copies); returns new (board, age, ko, player, status)."""
    n, size, board, age, ko, player = _canon(board, age, ko, player)
    board, age, ko, player = board.copy(), age.copy(), ko.copy(), player.copy()
    status = np.zeros(n, np.int32)
    lib().og_do_move_batch(n, size, board, age, ko, player, np.ascontiguousarray(moves, np.int32), status)
    return board.reshape(n, size, size), age.reshape(n, size, size), ko, player, status


def ladder_query(board, ko, player, action, prey=-1, escape=False):
    board = np.ascontiguousarray(board, dtype=np.int8)
    size = board.shape[-1]
    return bool(lib().og_ladder_query(size, board.reshape(-1), int(ko), int(player), int(action),
                                      int(prey), int(bool(escape))))
