"""rag_features_static: the feature entry point for hosts that are not allowed to allocate.

An engine that partitions its device memory once at start-up (or lives inside somebody else's
memory pool, or may not risk the device-wide stall of an allocation in its move loop) cannot hand
rag_features a workspace.  rag_features_static keeps that scratch inside the library instead.
This test drives it through the C ABI exactly as such a host would -- states and output buffer
are the caller's, nothing else is passed -- and checks, twice in a row on different batches, that
the planes are bit-for-bit those of the ordinary entry point (the second call runs on whatever
the first one left in the shared array).  The file sorts last so that a run stopped at the first
failure has already exercised everything else."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import go_oracle
from rocalphago_b200 import BatchedGameState, Preprocess, _cabi
from rocalphago_b200.state import _ptr, _stream

pytestmark = pytest.mark.gpu


def _both(batch, plane_src):
    """(planes from rag_features_static, planes from rag_features) for one batch."""
    L = _cabi.lib()
    n, S, F = len(batch), batch.size, len(plane_src)
    src = plane_src.ctypes.data_as(ctypes.c_void_p)
    got = torch.full((n, F, S, S), 7, dtype=torch.uint8, device=batch.device)
    _cabi.check(L.rag_features_static(_ptr(batch.packed), n, S, src, F, 0, 0, 64, _ptr(got), _stream()))
    want = torch.full((n, F, S, S), 9, dtype=torch.uint8, device=batch.device)
    ws = torch.empty(int(L.rag_features_workspace_bytes(n)), dtype=torch.uint8, device=batch.device)
    _cabi.check(L.rag_features(_ptr(batch.packed), n, S, src, F, 0, 0, 64, _ptr(want), _ptr(ws), ws.numel(), _stream()))
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_static_scratch_matches_workspace_entry_point_twice():
    L = _cabi.lib()
    limit = L.rag_features_static_max_positions()
    assert limit == 256
    pp = Preprocess()                                   # the 48 default planes, ladders included
    first = BatchedGameState.random_playouts(limit, seed=4242)
    second = BatchedGameState.random_playouts(77, seed=1717, max_ply=300, stride=4)
    for batch in (first, second):                       # the second call reuses the array as left by the first
        got, want = _both(batch, pp.plane_src)
        assert want.sum() > 0 and got.shape == want.shape
        assert np.array_equal(got, want)
    # a 9x9 batch right after the 19x19 ones, other planes, other order
    small = BatchedGameState.random_playouts(40, size=9, seed=5, max_ply=90, stride=3)
    got, want = _both(small, Preprocess(["ladder_escape", "legal", "liberties"]).plane_src)
    assert np.array_equal(got, want)


def test_static_scratch_against_the_oracle_and_without_touching_the_allocator():
    """The same entry point checked on its own: planes equal to the CPU oracle's on oracle-made positions,
    and the device's free memory the same before and after the calls (nothing is allocated or freed)."""
    L = _cabi.lib()
    b, a, ko, pl = go_oracle.random_playouts(200, size=19, seed=606, max_ply=380, stride=7)
    want = go_oracle.features(b, a, ko, pl, go_oracle.DEFAULT_FEATURES)
    batch = BatchedGameState.from_arrays(b, a, ko, pl)
    pp = Preprocess()
    src = pp.plane_src.ctypes.data_as(ctypes.c_void_p)
    got = torch.zeros((200, 48, 19, 19), dtype=torch.uint8, device="cuda")
    _cabi.check(L.rag_features_static(_ptr(batch.packed), 200, 19, src, 48, 0, 0, 64, _ptr(got), _stream()))   # set-up call
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        got.fill_(5)
        _cabi.check(L.rag_features_static(_ptr(batch.packed), 200, 19, src, 48, 0, 0, 64, _ptr(got), _stream()))
        torch.cuda.synchronize()
        assert np.array_equal(got.cpu().numpy(), want)
    assert torch.cuda.mem_get_info()[0] == free_before


def test_static_scratch_refuses_oversized_batches():
    L = _cabi.lib()
    n = L.rag_features_static_max_positions() + 1
    batch = BatchedGameState.random_playouts(n, seed=1)
    src = Preprocess(["board"]).plane_src
    out = torch.zeros((n, 3, 19, 19), dtype=torch.uint8, device="cuda")
    rc = L.rag_features_static(_ptr(batch.packed), n, 19, src.ctypes.data_as(ctypes.c_void_p), 3, 0, 0, 64,
                               _ptr(out), _stream())
    assert rc == -3 and b"at most 256" in L.rag_last_error()
    torch.cuda.synchronize()
    assert int(out.sum()) == 0                           # nothing was launched
