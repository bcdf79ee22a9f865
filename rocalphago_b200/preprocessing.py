"""Preprocess: GameState -> one-hot feature planes, computed on the GPU.

Drop-in for AlphaGo/preprocessing/preprocessing.py (names, feature order, shapes and error
text follow preprocessing.py:209-294).  `state_to_tensor(state)` keeps the reference
signature; `states_to_tensor` / `to_device` are the batched forms the reference lacks.
"""
import ctypes

import numpy as np
import torch

from . import _cabi
from .state import BatchedGameState, _ptr, _stream, _require_cuda

# name -> (number of planes, first canonical plane)   preprocessing.py:209-258
FEATURES = {
    "board": {"size": 3, "plane": 0},
    "ones": {"size": 1, "plane": 3},
    "turns_since": {"size": 8, "plane": 4},
    "liberties": {"size": 8, "plane": 12},
    "capture_size": {"size": 8, "plane": 20},
    "self_atari_size": {"size": 8, "plane": 28},
    "liberties_after": {"size": 8, "plane": 36},
    "ladder_capture": {"size": 1, "plane": 44},
    "ladder_escape": {"size": 1, "plane": 45},
    "sensibleness": {"size": 1, "plane": 46},
    "zeros": {"size": 1, "plane": 47},
    "legal": {"size": 1, "plane": 48},
    # not in the reference: value.py:16 takes 49 planes but preprocessing.py defines no 49th.
    # Here it is the player-colour plane of the AlphaGo paper: all ones iff BLACK is to move.
    "color": {"size": 1, "plane": 49},
}

# preprocessing.py:260-263
DEFAULT_FEATURES = [
    "board", "ones", "turns_since", "liberties", "capture_size",
    "self_atari_size", "liberties_after", "ladder_capture", "ladder_escape",
    "sensibleness", "zeros"]

LAYOUT_NCHW_U8, LAYOUT_NCHW_F32, LAYOUT_CONV_IN = 0, 1, 2
DTYPE_BF16, DTYPE_FP16, DTYPE_FP16X3, DTYPE_FP16C32, DTYPE_FP16C64 = 0, 1, 2, 3, 4        # include/rocalphago_b200.h: RAG_DTYPE_*
DTYPE_CODES = {"bf16": DTYPE_BF16, "fp16": DTYPE_FP16, "fp16x3": DTYPE_FP16X3, "fp16c32": DTYPE_FP16C32, "fp16c64": DTYPE_FP16C64}
_MOVE_PLANES = np.asarray([46, 48], np.int32)      # sensibleness, legal


_POINTS = {}


def _points(size):
    """[(x, y)] indexed by the flat point x*S+y (util.py:15-17), cached per board size."""
    if size not in _POINTS:
        _POINTS[size] = [(p // size, p % size) for p in range(size * size)]
    return _POINTS[size]


class _Workspace(object):
    """Grows-only scratch for the analyse step, one per (device)."""

    def __init__(self):
        self.buf = None

    def get(self, n, device):
        need = int(_cabi.lib().rag_features_workspace_bytes(n))
        if self.buf is None or self.buf.numel() < need or self.buf.device != device:
            self.buf = torch.empty(need, dtype=torch.uint8, device=device)
        return self.buf


def analyze_into(batch, ws, with_ladders=True):
    """rag_features_analyze -- or, for a batch built from states that enforce positional superko,
    rag_features_analyze_superko -- of `batch` into the workspace tensor `ws`, on the current stream."""
    L = _cabi.lib()
    with torch.cuda.device(batch.device):
        if getattr(batch, "superko", None) is not None:
            sk = batch.superko.struct()
            _cabi.check(L.rag_features_analyze_superko(_ptr(batch.packed), len(batch), batch.size, int(bool(with_ladders)),
                                                       ctypes.byref(sk), _ptr(ws), ws.numel(), _stream()))
        else:
            _cabi.check(L.rag_features_analyze(_ptr(batch.packed), len(batch), batch.size, int(bool(with_ladders)),
                                               _ptr(ws), ws.numel(), _stream()))


def check_workspace(ws):
    """rag_features_status: waits for the current stream and raises if a ladder kernel gave up on `ws`
    (its watchdog, include/rocalphago_b200.h).  Called where results are about to be read on the host."""
    with torch.cuda.device(ws.device):
        _cabi.check(_cabi.lib().rag_features_status(_ptr(ws), _stream()))


class Preprocess(object):
    """a class to convert from AlphaGo GameState objects to tensors of one-hot
    features for NN inputs (preprocessing.py:266-294)"""

    def __init__(self, feature_list=DEFAULT_FEATURES):
        self.output_dim = 0
        self.feature_list = feature_list
        self.processors = [None] * len(feature_list)
        src = []
        for i in range(len(feature_list)):
            feat = feature_list[i].lower()
            if feat in FEATURES:
                self.processors[i] = FEATURES[feat].get("function", feat)
                self.output_dim += FEATURES[feat]["size"]
                src += list(range(FEATURES[feat]["plane"], FEATURES[feat]["plane"] + FEATURES[feat]["size"]))
            else:
                raise ValueError("uknown feature: %s" % feat)        # sic, preprocessing.py:285
        self.plane_src = np.asarray(src, np.int32)
        self._ws = _Workspace()

    # ------------------------------------------------------------------ device-level
    def analyze(self, batch, with_ladders=None):
        """Runs the analysis + ladder kernels; returns the workspace tensor the emits read."""
        _require_cuda()
        if with_ladders is None:
            with_ladders = bool(np.any((self.plane_src == 44) | (self.plane_src == 45)))
        ws = self._ws.get(len(batch), batch.device)
        analyze_into(batch, ws, with_ladders)
        return ws

    def emit(self, ws, n, size, layout=LAYOUT_NCHW_U8, plane_src=None, conv_dtype=DTYPE_BF16,
             conv_channels=64, out=None):
        src = self.plane_src if plane_src is None else np.ascontiguousarray(plane_src, np.int32)
        F = len(src)
        dev = ws.device
        L = _cabi.lib()
        if out is None:
            if layout == LAYOUT_NCHW_U8:
                out = torch.empty((n, F, size, size), dtype=torch.uint8, device=dev)
            elif layout == LAYOUT_NCHW_F32:
                out = torch.empty((n, F, size, size), dtype=torch.float32, device=dev)
            else:
                rows = int(L.rag_conv_rows(n, size))
                out = torch.zeros((rows, conv_channels), device=dev,
                                  dtype=torch.bfloat16 if conv_dtype == DTYPE_BF16 else torch.float16)
        with torch.cuda.device(dev):
            _cabi.check(L.rag_features_emit(_ptr(ws), n, size, src.ctypes.data_as(ctypes.c_void_p), F, layout,
                                            conv_dtype, conv_channels, _ptr(out), _stream()))
        return out

    def to_device(self, batch, layout=LAYOUT_NCHW_U8, **kw):
        """Feature planes of a BatchedGameState as a CUDA tensor in `layout`."""
        ws = self.analyze(batch)
        return self.emit(ws, len(batch), batch.size, layout=layout, **kw)

    def legal_move_indices(self, ws, n, size):
        """get_legal_moves() of every position in reference order (go.py:465-480): legal non-eye
        points x-major, then legal own eyes x-major.  Returns (state index, flat point x*S+y) of
        every move, both int64 arrays sorted by state, and the number of moves per state."""
        P = size * size
        two = self.emit(ws, n, size, LAYOUT_NCHW_U8, plane_src=_MOVE_PLANES).cpu().numpy().reshape(n, 2, P)
        two[:, 1] &= 1 - two[:, 0]                       # own eyes only where not already listed
        which, col = np.nonzero(two.reshape(n, 2 * P))   # row-major: by state, non-eyes before eyes, then by point
        return which, col % P, np.bincount(which, minlength=n)

    def legal_move_lists(self, ws, n, size):
        """The same as lists of (x, y) tuples, one list per position."""
        _, flat, counts = self.legal_move_indices(ws, n, size)
        points = _points(size)
        moves = [points[p] for p in flat.tolist()]
        ends = np.cumsum(counts).tolist()
        return [moves[e - c:e] for e, c in zip(ends, counts.tolist())]

    # ------------------------------------------------------------------ reference-shaped API
    def states_to_tensor(self, states):
        """Batched state_to_tensor: (N, F, S, S) uint8 numpy array for a list of GameStates."""
        if len(states) == 0:
            return np.zeros((0, self.output_dim, 0, 0), np.uint8)
        batch = states if isinstance(states, BatchedGameState) else BatchedGameState.from_states(states)
        ws = self.analyze(batch)
        out = self.emit(ws, len(batch), batch.size, layout=LAYOUT_NCHW_U8)
        check_workspace(ws)
        return out.cpu().numpy()

    def state_to_tensor(self, state):
        """Convert a GameState to a Theano-compatible tensor: (1, F, S, S) float64 of 0/1
        (preprocessing.py:287-294)."""
        return self.states_to_tensor([state]).astype(np.float64)


# ---- the reference's module-level feature functions (preprocessing.py:14-205) ----------------------
#
# Each returns the float (k, S, S) planes of ONE feature for one state, as the reference's functions of
# the same names do; FEATURES[name]["function"] and Preprocess.processors point at them.  They run the
# same kernels as Preprocess (one position, one feature), so they are for compatibility, not speed.

_SINGLE = {}


def _feature(state, name):
    if name not in _SINGLE:
        _SINGLE[name] = Preprocess([name])
    return _SINGLE[name].state_to_tensor(state)[0].astype(np.float64)


def _capped(state, name, maximum):
    """One-hot planes saturating at `maximum` instead of 8: plane min(v, maximum - 1).  The kernels
    saturate at 8 planes, so 1 <= maximum <= 8 can be folded from them; more cannot."""
    if not 1 <= maximum <= 8:
        raise ValueError("maximum must be in [1, 8] on this path (got %d)" % maximum)
    planes = _feature(state, name)
    out = planes[:maximum].copy()
    out[maximum - 1] = planes[maximum - 1:].max(axis=0)
    return out


def get_board(state):
    """own stones, opponent stones, empty points (preprocessing.py:14-22)"""
    return _feature(state, "board")


def get_turns_since(state, maximum=8):
    """one-hot age of every stone, saturating (preprocessing.py:25-37)"""
    return _capped(state, "turns_since", maximum)


def get_liberties(state, maximum=8):
    """one-hot liberty count of every stone's group (preprocessing.py:40-56)"""
    return _capped(state, "liberties", maximum)


def get_capture_size(state, maximum=8):
    """stones captured by playing at each legal point (preprocessing.py:59-86)"""
    return _capped(state, "capture_size", maximum)


def get_self_atari_size(state, maximum=8):
    """size of the own group put into atari by playing at each legal point (preprocessing.py:89-126)"""
    return _capped(state, "self_atari_size", maximum)


def get_liberties_after(state, maximum=8):
    """liberties of the played stone's group after each legal move (preprocessing.py:129-169)"""
    return _capped(state, "liberties_after", maximum)


def get_ladder_capture(state):
    """legal moves that capture in a ladder (preprocessing.py:172-178)"""
    return _feature(state, "ladder_capture")


def get_ladder_escape(state):
    """legal moves that escape a ladder (preprocessing.py:181-187)"""
    return _feature(state, "ladder_escape")


def get_sensibleness(state):
    """legal moves that do not fill an own eye (preprocessing.py:190-196)"""
    return _feature(state, "sensibleness")


def get_legal(state):
    """legal moves, own eyes included (preprocessing.py:199-205)"""
    return _feature(state, "legal")


for _name, _fn in (("board", get_board), ("turns_since", get_turns_since), ("liberties", get_liberties),
                   ("capture_size", get_capture_size), ("self_atari_size", get_self_atari_size),
                   ("liberties_after", get_liberties_after), ("ladder_capture", get_ladder_capture),
                   ("ladder_escape", get_ladder_escape), ("sensibleness", get_sensibleness), ("legal", get_legal),
                   ("ones", lambda state: np.ones((1, state.size, state.size))),
                   ("zeros", lambda state: np.zeros((1, state.size, state.size))),
                   ("color", lambda state: _feature(state, "color"))):
    FEATURES[_name]["function"] = _fn
