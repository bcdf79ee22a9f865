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


# ---- superko (go.py:242-265): states created with enforce_superko=True ---------------------------

def zobrist_table(size):
    """hash_lookup of go.py:58-62 as one array [2][size*size]: [0] BLACK, [1] WHITE.  The reference
    draws WHITE's table first from RandomState(0)."""
    rng = np.random.RandomState(0)
    white = rng.randint(np.iinfo(np.uint64).max, size=(size, size), dtype="uint64")
    black = rng.randint(np.iinfo(np.uint64).max, size=(size, size), dtype="uint64")
    return np.ascontiguousarray(np.stack([black.reshape(-1), white.reshape(-1)]))


def features_superko(board, age, ko, player, sk, feature_list=DEFAULT_FEATURES, with_moves=False):
    """As `features`, for states that enforce positional superko.  sk: dict with enforce (n,),
    has_handicap (n,), hist_len (n,), hash (n,) uint64, prev (flat uint64), prev_off (n+1,),
    masks (n, 3, S*S) uint8 = points in history[0::2] / history[1::2] / handicaps."""
    n, size, board, age, ko, player = _canon(board, age, ko, player)
    ids = np.asarray([FEATURE_IDS[f.lower()] for f in feature_list], dtype=np.int32)
    F = int(sum(lib().og_feature_planes(int(i)) for i in ids))
    out = np.empty((n, F, size, size), dtype=np.uint8)
    L = lib()
    if not getattr(L, "_sk_ready", False):
        u64p = np.ctypeslib.ndpointer(np.uint64, flags="C")
        i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
        i16p = np.ctypeslib.ndpointer(np.int16, flags="C")
        i8p = np.ctypeslib.ndpointer(np.int8, flags="C")
        u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
        L.og_features_superko.restype = ctypes.c_int
        L.og_features_superko.argtypes = [ctypes.c_int, ctypes.c_int, i8p, i16p, i32p, i32p, i32p, ctypes.c_int, u8p,
                                          u64p, i32p, i32p, i32p, u64p, u64p, i32p, u8p, ctypes.c_void_p, ctypes.c_void_p]
        L._sk_ready = True
    c = lambda a, t: np.ascontiguousarray(a, dtype=t)  # noqa: E731
    prev = c(sk["prev"], np.uint64)
    if prev.size == 0:
        prev = np.zeros(1, np.uint64)
    moves = np.full((n, size * size), -1, np.int16)
    counts = np.zeros((n, 2), np.int32)
    got = L.og_features_superko(n, size, board, age, ko, player, ids, len(ids), out.reshape(-1), zobrist_table(size).reshape(-1),
                                c(sk["enforce"], np.int32), c(sk["has_handicap"], np.int32), c(sk["hist_len"], np.int32),
                                c(sk["hash"], np.uint64), prev, c(sk["prev_off"], np.int32),
                                c(sk["masks"], np.uint8).reshape(-1), moves.ctypes.data_as(ctypes.c_void_p),
                                counts.ctypes.data_as(ctypes.c_void_p))
    assert got == F, got
    if with_moves:
        return out, [moves[i, :counts[i, 0]].astype(np.int64) for i in range(n)], counts[:, 1].copy()
    return out


def winners(board, age, ko, player, komi=7.5):
    """GameState.get_winner() (go.py:482-506) of each position, no passes made: +1 BLACK, -1 WHITE, 0 tie."""
    res = rollout_winners(board, age, ko, player, limit=0, seed=0, komi=komi)
    return res * np.asarray(player, np.int32).reshape(-1)                                   # rollout results are relative to the side to move


def rollout_winners(board, age, ko, player, limit=500, seed=0, komi=7.5):
    """+1 / -1 / 0 for the player to move after a uniformly random sensible-move rollout (see go_oracle.c)."""
    n, size, board, age, ko, player = _canon(board, age, ko, player)
    L = lib()
    L.og_rollout_winners.restype = None
    L.og_rollout_winners.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_void_p, ctypes.c_int, ctypes.c_uint64, ctypes.c_double, ctypes.c_void_p]
    out = np.zeros(n, np.int32)
    L.og_rollout_winners(n, size, board.ctypes.data, age.ctypes.data, ko.ctypes.data, player.ctypes.data,
                         int(limit), int(seed), float(komi), out.ctypes.data)
    return out
