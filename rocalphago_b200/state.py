"""Packed, batched game states in HBM.

The device twin of the `GameState` fields the feature planes depend on
(AlphaGo/go.py:18-22,55: board, current_player, ko, stone_ages).  Reference-style
`GameState` objects are accepted anywhere a batch is built; they are read, never modified.
"""
import ctypes

import numpy as np
import torch

from . import _cabi

STATE_WORDS = 192
PASS_MOVE = None


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _np_ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _require_cuda():
    if not torch.cuda.is_available():
        raise _cabi.RagError("rocalphago_b200 needs a CUDA device (sm_100a); there is no CPU path")


class PositionView(object):
    """The fields of one position as plain arrays: a minimal GameState look-alike
    (.size .board .stone_ages .ko .current_player) for code that only reads them."""

    def __init__(self, board, stone_ages, ko, current_player):
        self.board = np.asarray(board)
        self.size = int(self.board.shape[0])
        self.stone_ages = np.asarray(stone_ages)
        self.ko = None if ko is None or (np.isscalar(ko) and ko < 0) else (
            tuple(ko) if not np.isscalar(ko) else (int(ko) // self.size, int(ko) % self.size))
        self.current_player = int(current_player)


def arrays_from_states(states):
    """(board int8 (n,S,S), age int16 (n,S,S), ko int32 (n,), player int32 (n,)) from GameState-likes."""
    n = len(states)
    size = states[0].size
    if not all(st.size == size for st in states):
        raise ValueError("all states must have the same size")       # policy.py:39-40
    board = np.stack([st.board for st in states]).astype(np.int8, copy=False)
    age = np.stack([st.stone_ages for st in states]).astype(np.int16, copy=False)
    ko = np.array([-1 if st.ko is None else st.ko[0] * size + st.ko[1] for st in states], np.int32)
    player = np.array([st.current_player for st in states], np.int32)
    return board, age, ko, player


def pack_arrays(board, age, ko, player, pinned=False):
    """Host packing (rag_pack_states) -> torch int32 (n, 192) on the host."""
    board = np.ascontiguousarray(board, np.int8)
    n, size = board.shape[0], board.shape[-1]
    age = np.ascontiguousarray(age, np.int16)
    ko = np.ascontiguousarray(ko, np.int32)
    player = np.ascontiguousarray(player, np.int32)
    packed = torch.empty((n, STATE_WORDS), dtype=torch.int32, pin_memory=bool(pinned and torch.cuda.is_available()))
    _cabi.check(_cabi.lib().rag_pack_states(n, size, _np_ptr(board), _np_ptr(age), _np_ptr(ko), _np_ptr(player),
                                            ctypes.c_void_p(packed.data_ptr())))
    return packed


class BatchedGameState(object):
    """n positions of one board size, packed (n, 192) int32 on a CUDA device."""

    def __init__(self, packed, size):
        self.packed = packed
        self.size = int(size)

    def __len__(self):
        return int(self.packed.shape[0])

    @property
    def device(self):
        return self.packed.device

    @classmethod
    def from_arrays(cls, board, age, ko, player, device=None):
        _require_cuda()
        host = pack_arrays(board, age, ko, player, pinned=len(ko) >= 1024)      # pinning costs more than it saves on small batches
        size = np.asarray(board).shape[-1]
        return cls(host.to(device or "cuda", non_blocking=True), size)

    @classmethod
    def from_states(cls, states, device=None):
        return cls.from_arrays(*arrays_from_states(states), device=device)

    @classmethod
    def random_playouts(cls, n, size=19, seed=20261106, max_ply=400, stride=8, device=None, with_log=False):
        """n positions from seeded random playouts run on the device."""
        _require_cuda()
        dev = torch.device(device or "cuda")
        with torch.cuda.device(dev):
            packed = torch.zeros((n, STATE_WORDS), dtype=torch.int32, device=dev)
            L = _cabi.lib()
            log, log_stride = None, 0
            if with_log:
                quota = L.rag_playout_quota(max_ply, stride)
                warps = (n + quota - 1) // quota
                log_stride = 4 * (max_ply + 2) + 8
                log = torch.full((warps, log_stride), -3, dtype=torch.int32, device=dev)
            _cabi.check(L.rag_random_playouts(_ptr(packed), n, size, ctypes.c_uint64(seed), max_ply, stride,
                                              _ptr(log) if log is not None else None, log_stride, _stream()))
        out = cls(packed, size)
        return (out, log) if with_log else out

    def do_move(self, moves):
        """GameState.do_move for every position (flat index x*S+y, -1 or None = pass), in place.
        Returns an int32 tensor: 0 played, -1 illegal (that position is left untouched)."""
        mv = [-1 if m is None else (m if np.isscalar(m) else m[0] * self.size + m[1]) for m in moves] \
            if not torch.is_tensor(moves) else moves
        mv = torch.as_tensor(mv, dtype=torch.int32).to(self.device).contiguous()
        status = torch.empty(len(self), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().rag_do_move(_ptr(self.packed), len(self), self.size, _ptr(mv), _ptr(status),
                                                _stream()))
        return status

    def to_arrays(self):
        """(board, age (saturated at 7), ko, player) numpy arrays."""
        host = self.packed.cpu().contiguous()
        n, S = len(self), self.size
        board = np.empty((n, S, S), np.int8)
        age = np.empty((n, S, S), np.int16)
        ko = np.empty(n, np.int32)
        player = np.empty(n, np.int32)
        _cabi.check(_cabi.lib().rag_unpack_states(n, S, ctypes.c_void_p(host.data_ptr()), _np_ptr(board),
                                                  _np_ptr(age), _np_ptr(ko), _np_ptr(player)))
        return board, age, ko, player

    def views(self):
        board, age, ko, player = self.to_arrays()
        return [PositionView(board[i], age[i], int(ko[i]), int(player[i])) for i in range(len(self))]
