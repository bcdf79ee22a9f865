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
AGELESS_WORD = 5 * 32 + 8          # meta plane words 8..26 (include/rocalphago_b200.h)
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
    # one preallocated array per field, filled state by state (numpy converts float64 boards / int64 ages
    # on assignment): a third of the time of np.stack + astype at 4,096 states
    board = np.empty((n, size, size), np.int8)
    age = np.empty((n, size, size), np.int16)
    for i, st in enumerate(states):
        board[i] = st.board
        age[i] = st.stone_ages
    ko = np.fromiter((-1 if st.ko is None else st.ko[0] * size + st.ko[1] for st in states), np.int32, n)
    player = np.fromiter((st.current_player for st in states), np.int32, n)
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


_ZOBRIST = {}


def zobrist_table(size):
    """GameState.hash_lookup (go.py:58-62) laid out for the device: uint64 [2][1024] indexed by
    q = x*32 + y, [0] BLACK, [1] WHITE.  The reference draws WHITE's table first from RandomState(0)."""
    if size not in _ZOBRIST:
        rng = np.random.RandomState(0)
        white = rng.randint(np.iinfo(np.uint64).max, size=(size, size), dtype="uint64")
        black = rng.randint(np.iinfo(np.uint64).max, size=(size, size), dtype="uint64")
        t = np.zeros((2, 32, 32), np.uint64)
        t[0, :size, :size], t[1, :size, :size] = black, white
        _ZOBRIST[size] = t.reshape(2, 1024)
    return _ZOBRIST[size]


class SuperkoData(object):
    """What is_positional_superko (go.py:242-265) reads from each state, on the device (rag_superko).
    Built by BatchedGameState.from_states when a state has enforce_superko set."""

    def __init__(self, states, size, device):
        n = len(states)
        masks = np.zeros((n, 3, 32), np.uint32)
        enforce, handi, hist_len = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.int32)
        hashes = np.zeros(n, np.uint64)
        prev, off = [], np.zeros(n + 1, np.int32)
        for i, st in enumerate(states):
            if getattr(st, "enforce_superko", False):
                for k, mv in enumerate(st.history):
                    if mv is not None:
                        masks[i, k & 1, mv[0]] |= np.uint32(1 << mv[1])
                for mv in st.handicaps:
                    masks[i, 2, mv[0]] |= np.uint32(1 << mv[1])
                enforce[i], handi[i], hist_len[i] = 1, int(len(st.handicaps) > 0), len(st.history)
                hashes[i] = np.uint64(st.current_hash)
                prev.append(np.fromiter((int(h) for h in st.previous_hashes), np.uint64, len(st.previous_hashes)))
                off[i + 1] = off[i] + len(prev[-1])
            else:
                off[i + 1] = off[i]
        self._upload(size, device, enforce, handi, hist_len, hashes,
                     np.concatenate(prev) if prev else np.zeros(0, np.uint64), off, masks)

    @classmethod
    def from_arrays(cls, size, device, enforce, has_handicap, hist_len, hashes, prev, prev_off, point_masks):
        """From plain arrays: point_masks uint8 (n, 3, S*S) = points of history[0::2] / history[1::2] /
        handicaps (flat index x*S+y); prev = all previous_hashes, position b's in prev_off[b]:prev_off[b+1]."""
        self = cls.__new__(cls)
        pm = np.asarray(point_masks).reshape(len(enforce), 3, size, size).astype(bool)
        masks = np.zeros((len(enforce), 3, 32), np.uint32)
        masks[:, :, :size] = (pm * (np.uint32(1) << np.arange(size, dtype=np.uint32))).sum(axis=3).astype(np.uint32)
        self._upload(size, device, np.asarray(enforce, np.int32), np.asarray(has_handicap, np.int32),
                     np.asarray(hist_len, np.int32), np.asarray(hashes, np.uint64), np.asarray(prev, np.uint64),
                     np.asarray(prev_off, np.int32), masks)
        return self

    def _upload(self, size, device, enforce, has_handicap, hist_len, hashes, prev, prev_off, masks):
        masks = masks * enforce[:, None, None].astype(np.uint32)          # a position that does not enforce superko has no history here
        flags = (enforce | (has_handicap << 1) | (hist_len << 8)).astype(np.int32) * enforce      # 0 = evaluated without superko
        if prev.size == 0:
            prev = np.zeros(1, np.uint64)
        as_i64 = lambda a: torch.as_tensor(np.ascontiguousarray(a).view(np.int64)).to(device)  # noqa: E731  (only the bits travel)
        self.zobrist = as_i64(zobrist_table(size))
        self.hash, self.prev = as_i64(hashes), as_i64(prev)
        self.prev_off = torch.as_tensor(np.ascontiguousarray(prev_off)).to(device)
        self.masks = torch.as_tensor(np.ascontiguousarray(masks).view(np.int32)).to(device)
        self.flags = torch.as_tensor(flags).to(device)

    def struct(self):
        """ctypes rag_superko pointing at the tensors (keep `self` alive while it is in use)."""
        return _cabi.RagSuperko(*[ctypes.c_void_p(t.data_ptr()) for t in
                                  (self.zobrist, self.hash, self.prev, self.prev_off, self.masks, self.flags)])


class BatchedGameState(object):
    """n positions of one board size, packed (n, 192) int32 on a CUDA device.  `superko` is None unless
    the batch was built from states that enforce positional superko (then features, move lists and
    network evaluations honour it; states advanced on the device -- do_move, playouts -- carry no
    history and never have it)."""

    def __init__(self, packed, size, superko=None):
        self.packed = packed
        self.size = int(size)
        self.superko = superko

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
        batch = cls.from_arrays(*arrays_from_states(states), device=device)
        if any(getattr(st, "enforce_superko", False) for st in states):
            batch.superko = SuperkoData(states, batch.size, batch.device)
        return batch

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
        self.superko = None                 # the device keeps no move history: from here on superko is not enforced
        with torch.cuda.device(self.device):
            _cabi.check(_cabi.lib().rag_do_move(_ptr(self.packed), len(self), self.size, _ptr(mv), _ptr(status),
                                                _stream()))
        return status

    def forget_ages(self):
        """What GameState.copy() does to stone_ages (go.py:186-217 does not copy them): every stone on
        the board loses its age -- it is in no turns_since plane and later moves do not age it -- while
        stones played from here on age as usual.  In place; returns self."""
        stones = self.packed[:, 0:32] | self.packed[:, 32:64]
        self.packed[:, 64:160] = 0
        self.packed[:, AGELESS_WORD:AGELESS_WORD + 24] = stones[:, :24]
        return self

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
