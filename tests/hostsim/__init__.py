"""Host stand-in for one warp running rocalphago_b200/csrc/go_bb.cuh -- TEST INFRASTRUCTURE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libhostsim.so")
_SRC = [os.path.join(_HERE, "hostsim.cpp"),
        os.path.join(_HERE, "..", "..", "rocalphago_b200", "csrc", "go_bb.cuh")]
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in _SRC):
            subprocess.check_call(["g++", "-O2", "-fPIC", "-shared", "-x", "c++", "-o", _SO, _SRC[0]])
        L = ctypes.CDLL(_SO)
        i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
        L.hostsim_features.restype = ctypes.c_int
        L.hostsim_features.argtypes = [ctypes.c_int, ctypes.c_int, np.ctypeslib.ndpointer(np.int8, flags="C"),
                                       np.ctypeslib.ndpointer(np.int16, flags="C"), i32p, i32p,
                                       np.ctypeslib.ndpointer(np.uint8, flags="C"), ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_int64]
        _lib = L
    return _lib


def features49(board, age, ko, player, item_cap=0):
    """uint8 (n, 49, S, S): the 48 default planes + legal, computed by the lane-emulated device code."""
    board = np.ascontiguousarray(board, np.int8)
    n, S = board.shape[0], board.shape[-1]
    out = np.empty((n, 49, S, S), np.uint8)
    stats = np.zeros(16, np.int64)
    items = np.zeros(item_cap, np.int64) if item_cap else None
    lib().hostsim_features(n, S, board.reshape(n, -1), np.ascontiguousarray(age, np.int16).reshape(n, -1),
                           np.ascontiguousarray(ko, np.int32), np.ascontiguousarray(player, np.int32),
                           out.reshape(-1), stats.ctypes.data_as(ctypes.c_void_p),
                           items.ctypes.data_as(ctypes.c_void_p) if item_cap else None, item_cap)
    if item_cap:
        return out, stats, items[:min(item_cap, int(stats[3]))]
    return out, stats


def features49_superko(board, age, ko, player, sk, zobrist):
    """The same for states that enforce positional superko; `sk` as oracle.go_oracle.features_superko
    takes it, `zobrist` = go_oracle.zobrist_table(size)."""
    board = np.ascontiguousarray(board, np.int8)
    n, S = board.shape[0], board.shape[-1]
    out = np.empty((n, 49, S, S), np.uint8)
    L = lib()
    vp = ctypes.c_void_p
    L.hostsim_features_superko.restype = ctypes.c_int
    L.hostsim_features_superko.argtypes = [ctypes.c_int, ctypes.c_int] + [vp] * 12
    keep = [board.reshape(n, -1), np.ascontiguousarray(age, np.int16).reshape(n, -1), np.ascontiguousarray(ko, np.int32),
            np.ascontiguousarray(player, np.int32), out, np.ascontiguousarray(zobrist, np.uint64),
            np.ascontiguousarray(sk["has_handicap"], np.int32), np.ascontiguousarray(sk["hist_len"], np.int32),
            np.ascontiguousarray(sk["hash"], np.uint64),
            np.ascontiguousarray(sk["prev"] if len(sk["prev"]) else np.zeros(1), np.uint64),
            np.ascontiguousarray(sk["prev_off"], np.int32), np.ascontiguousarray(sk["masks"], np.uint8)]
    L.hostsim_features_superko(n, S, *[a.ctypes.data_as(vp) for a in keep])
    return out
