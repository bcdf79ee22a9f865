"""BASELINE-size runs checked through size-independent properties (the oracle would take minutes
to hours at these sizes): determinism, plane algebra that holds for every legal Go position, the
do_move / feature consistency, and pipeline == single-stream results at batch 4096."""
import numpy as np
import pytest
import torch

from rocalphago_b200 import BatchedGameState, Preprocess
from rocalphago_b200 import preprocessing as P
from rocalphago_b200.models import CNNPolicy

pytestmark = pytest.mark.gpu

ALL = P.DEFAULT_FEATURES + ["legal", "color"]


def _planes(batch):
    return Preprocess(ALL).to_device(batch, P.LAYOUT_NCHW_U8)


def test_feature_sweep_262144_positions_invariants():
    """A quarter of BASELINE config 4's million positions in one call (the sweep itself is
    `bench.py --workload features`): plane algebra that every legal position satisfies."""
    n = 262144
    batch = BatchedGameState.random_playouts(n, seed=99)
    f = _planes(batch)                                   # (n, 50, 19, 19) uint8 on the device
    assert f.shape == (n, 50, 19, 19) and int(f.max()) == 1
    own, opp, empty = f[:, 0], f[:, 1], f[:, 2]
    assert torch.equal(own + opp + empty, torch.ones_like(own))                 # board planes partition the points
    assert torch.equal(f[:, 3], torch.ones_like(own)) and int(f[:, 47].sum()) == 0   # ones / zeros
    stones = own + opp
    assert torch.equal(f[:, 4:12].sum(1), stones)                               # turns_since: one-hot on stones
    assert torch.equal(f[:, 12:20].sum(1), stones)                              # liberties: one-hot on stones
    legal, sens = f[:, 49 - 1], f[:, 46]
    assert int((legal * stones).sum()) == 0 and int((sens > legal).sum()) == 0  # legal on empties; sensible within legal
    assert torch.equal(f[:, 20:28].sum(1), legal)                               # capture_size: one-hot on legal moves
    assert torch.equal(f[:, 36:44].sum(1), legal)                               # liberties_after: one-hot on legal moves
    assert int((f[:, 28:36].sum(1) > legal).sum()) == 0                         # self_atari within legal
    assert torch.equal(f[:, 28:36].sum(1), f[:, 36])                            # self-atari <=> exactly one liberty after
    assert int((f[:, 44] > legal).sum()) == 0 and int((f[:, 45] > legal).sum()) == 0   # ladder planes within legal
    color = f[:, 49]
    assert bool(((color.flatten(1).sum(1) == 0) | (color.flatten(1).sum(1) == 361)).all())
    # determinism: the same call again gives the same bytes (the ladder reader hands work around
    # between warps in whatever order they become idle)
    assert torch.equal(f, _planes(batch))
    # all game phases are present
    cnt = stones.flatten(1).sum(1)
    assert int(cnt.min()) == 0 and int(cnt.max()) > 250


def test_do_move_and_features_agree_on_131072_positions():
    """Playing a sensible move never fails; playing a point the legal plane excludes always does;
    afterwards the stone is there as `opponent` (the side to move flipped) with age 0."""
    n = 131072
    batch = BatchedGameState.random_playouts(n, seed=7, max_ply=300, stride=6)
    f = Preprocess(["sensibleness", "legal", "board"]).to_device(batch, P.LAYOUT_NCHW_U8)
    sens, legal = f[:, 0].flatten(1), f[:, 1].flatten(1)
    has_move = sens.sum(1) > 0
    first = sens.float().argmax(1).int()                                         # first sensible point
    moves = torch.where(has_move, first, torch.full_like(first, -1))
    after = BatchedGameState(batch.packed.clone(), 19)
    status = after.do_move(moves)
    assert int((status != 0).sum()) == 0
    g = Preprocess(["board", "turns_since"]).to_device(after, P.LAYOUT_NCHW_U8)
    idx = torch.arange(n, device=f.device)[has_move]
    mv = first[has_move].long()
    assert bool((g[idx, 1].flatten(1)[torch.arange(len(idx)), mv] == 1).all())   # now the opponent's stone
    assert bool((g[idx, 3].flatten(1)[torch.arange(len(idx)), mv] == 1).all())   # turns_since plane 0
    # an illegal empty point (suicide or ko), where one exists, is refused and leaves the state alone
    empty = f[:, 4].flatten(1)
    bad = (empty == 1) & (legal == 0)
    has_bad = bad.sum(1) > 0
    badmove = torch.where(has_bad, bad.float().argmax(1).int(), torch.zeros_like(first))
    again = BatchedGameState(batch.packed.clone(), 19)
    st = again.do_move(badmove)
    assert bool((st[has_bad] == -1).all()) and int(has_bad.sum()) > 100
    assert torch.equal(again.packed[has_bad], batch.packed[has_bad])


def test_policy_batch_4096_pipeline_equals_single_stream():
    """BASELINE's flagship shape (13x192, batch 4096): the two-stream pipeline returns bit-identical
    probabilities to the single-stream fused call, rows sum to one, and they depend on the position."""
    from rocalphago_b200.pipeline import EvalPipeline
    n = 4096
    net = CNNPolicy(P.DEFAULT_FEATURES, layers=12, filters_per_layer=192, dtype="fp16", seed=3)
    batches = [BatchedGameState.random_playouts(n, seed=40 + i) for i in range(3)]
    want = [net.model.eval_batch_device(b, net.preprocessor.plane_src).clone() for b in batches]
    pipe = EvalPipeline(net, n)
    got = []
    for i, b in enumerate(batches):
        t = pipe.submit(b.packed)
        got.append(pipe.result(t).clone())
    pipe.drain()
    for w, g in zip(want, got):
        assert torch.equal(w, g)
        assert torch.allclose(g.sum(1), torch.ones(n, device=g.device), atol=1e-5)
    assert not torch.equal(got[0], got[1])
