"""How the hot path spreads over the GPUs of one box: no data-path collective.

The independent unit of both kernel families is (request, kv-head): a GLA state with its q-group,
or a request's KV pages with its q-group.  Heads are sharded first, exactly as vLLM tensor
parallelism shards ``QKVParallelLinear`` upstream of the path (reference
vllm_hymeta/models/modeling_hymeta.py:401-430, 694-732: ``num_heads % tp == 0`` and kv-heads
divisible), up to the number of kv-heads; beyond that requests are sharded (28/4 heads admit
head-sharding 1, 2 or 4 ways only -- SURVEY F8).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List


@dataclass(frozen=True)
class ShardPlan:
    world: int
    rank: int
    tp: int            # head shards
    dp: int            # request shards
    tp_rank: int
    dp_rank: int
    q_heads: int       # rank-local q heads
    kv_heads: int      # rank-local kv heads
    q_head_start: int  # first global q head of this rank
    kv_head_start: int

    def requests(self, n_requests: int) -> range:
        """Contiguous slice of the global request list owned by this rank's request shard."""
        per = (n_requests + self.dp - 1) // self.dp
        lo = min(n_requests, self.dp_rank * per)
        return range(lo, min(n_requests, lo + per))


def plan(world: int, rank: int, num_heads: int = 28, num_kv_heads: int = 4) -> ShardPlan:
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad world/rank {world}/{rank}")
    tp = 1
    while tp * 2 <= world and num_kv_heads % (tp * 2) == 0 and num_heads % (tp * 2) == 0 \
            and world % (tp * 2) == 0:
        tp *= 2
    dp = world // tp
    tp_rank, dp_rank = rank % tp, rank // tp
    qh, kvh = num_heads // tp, num_kv_heads // tp
    return ShardPlan(world, rank, tp, dp, tp_rank, dp_rank, qh, kvh, tp_rank * qh, tp_rank * kvh)


def all_plans(world: int, num_heads: int = 28, num_kv_heads: int = 4) -> List[ShardPlan]:
    return [plan(world, r, num_heads, num_kv_heads) for r in range(world)]
