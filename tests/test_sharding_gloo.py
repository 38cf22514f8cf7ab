"""N > 1 host logic on CPU: world_size-2 (and 4) gloo process groups exercise the shard plan
bench.py uses and its max-over-ranks timing reduction.  No GPU, no kernels."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vllm_hymeta_b200 import sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_requests):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        p = sharding.plan(world, rank)
        # every rank publishes the (request, kv-head) units it owns; together they must tile the
        # whole problem exactly once
        owned = torch.zeros(n_requests, 4, dtype=torch.int32)
        reqs = p.requests(n_requests)
        owned[reqs.start:reqs.stop, p.kv_head_start:p.kv_head_start + p.kv_heads] = 1
        dist.all_reduce(owned, op=dist.ReduceOp.SUM)
        assert torch.equal(owned, torch.ones_like(owned)), owned
        # q heads follow their kv head (GQA group of 7)
        assert p.q_head_start == p.kv_head_start * 7 and p.q_heads == p.kv_heads * 7
        # timing reduction: the job time is the slowest rank's
        t = torch.tensor([1.0 + rank])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert t.item() == float(world)
        # whole-job token count
        tok = torch.tensor([len(reqs) * 8704 * p.kv_heads], dtype=torch.int64)
        dist.all_reduce(tok)
        assert tok.item() == n_requests * 8704 * 4
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_requests", [(2, 128), (4, 256)])
def test_shard_plan_over_gloo(world, n_requests):
    mp.spawn(_worker, args=(world, _free_port(), n_requests), nprocs=world, join=True)


def test_plan_geometry():
    for world, tp, dp in [(1, 1, 1), (2, 2, 1), (4, 4, 1), (8, 4, 2)]:
        ps = sharding.all_plans(world)
        assert all(p.tp == tp and p.dp == dp for p in ps)
        assert sorted((p.dp_rank, p.tp_rank) for p in ps) == [(d, t) for d in range(dp) for t in range(tp)]
    p = sharding.plan(8, 5)                     # second head shard of the second request shard
    assert (p.tp_rank, p.dp_rank, p.q_heads, p.kv_heads, p.q_head_start) == (1, 1, 7, 1, 7)
    assert list(p.requests(10)) == [5, 6, 7, 8, 9]
    # a synthetic 64/8-head geometry shards heads 8 ways
    assert sharding.plan(8, 3, 64, 8).tp == 8
    with pytest.raises(ValueError):
        sharding.plan(2, 2)
