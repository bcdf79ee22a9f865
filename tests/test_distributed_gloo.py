"""Host logic of the multi-GPU path on CPU: world_size-2 (and 3) gloo process groups check the
sharding and the result gather that bench.py / a serving process use at N > 1.  The evaluation
function is a stand-in (the real one needs a GPU); what is tested is that every position is
evaluated exactly once, by the right rank, and comes back in order."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rocalphago_b200 import distributed as D


def test_shard_bounds_cover_everything_once():
    for n in (0, 1, 7, 8, 4096, 4099):
        for world in (1, 2, 3, 8):
            cuts = [D.shard_bounds(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[i][1] == cuts[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in cuts]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
            assert sizes == D.shard_sizes(n, world)
    with pytest.raises(ValueError):
        D.shard_bounds(10, 2, 2)


def _worker(rank, world, port, n, tmp):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        packed = torch.arange(n * 4, dtype=torch.int32).reshape(n, 4)          # stands for packed states
        seen = []

        def fake_eval(sl):                                                       # "probabilities": a function of the state only
            seen.append(sl.clone())
            return torch.stack([sl[:, 0].float() * 0.5, sl[:, 3].float() + 1000 * rank], dim=1)

        out = D.evaluate_sharded(fake_eval, packed)
        b, e = D.shard_bounds(n, rank, world)
        assert len(seen) == 1 and torch.equal(seen[0], packed[b:e])             # this rank touched only its slice
        assert out.shape == (n, 2)
        owner = torch.zeros(n)
        for r in range(world):
            rb, re = D.shard_bounds(n, r, world)
            owner[rb:re] = r
        assert torch.equal(out[:, 0], packed[:, 0].float() * 0.5)
        assert torch.equal(out[:, 1], packed[:, 3].float() + 1000 * owner)      # each row came from its owner
        # the max-over-ranks reduction bench.py uses for its timing
        t = torch.tensor([10.0 + rank], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert float(t) == 10.0 + world - 1
        np.save(os.path.join(tmp, "ok_%d.npy" % rank), out.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 10), (2, 7), (3, 8)])
def test_sharded_evaluation_and_gather_over_gloo(world, n, tmp_path):
    port = 29500 + (os.getpid() % 1000) + world * 7 + n
    mp.spawn(_worker, args=(world, port, n, str(tmp_path)), nprocs=world, join=True)
    outs = [np.load(str(tmp_path / ("ok_%d.npy" % r))) for r in range(world)]
    for o in outs[1:]:
        assert np.array_equal(o, outs[0])                                        # every rank got the same full result
