#!/usr/bin/env python3
"""BASELINE.json configs 3 and 5 at N GPUs (one process per GPU; `python tools/bench_configs.py` for
N=1 or `python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1
tools/bench_configs.py` for N>1).  Device-event timing, barrier on both sides, max over ranks.

config 3: ONE 70B-shaped hybrid layer, prompts of 32 768 tokens with the 128-token meta
  prefix, head-sharded: gate pre-processing -> GLA chunked prefill (meta state init) ;
  fused RoPE + paged-KV write -> SWA+meta prefill (window 4095) ; branch merge with its ONE 8*T-byte
  all-reduce (the only exchange next to the path).  28/4 heads shard 1/2/4 ways; at 8 GPUs the plan
  is 4 head shards x 2 request shards.  `--prompts` 32K prompts per request shard: the
  job is fixed while heads are sharded (strong scaling to N=4) and doubles with the request shards.
config 5: GLA chunked-scan prefill only, 8 sequences x T for T = 2K..128K, (request, kv-head)
  chains spread over the GPUs (heads <= 4-way, then requests); strong scaling of a fixed job.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from vllm_hymeta_b200 import ops, sharding

D, H, HKV, M, BS = 128, 28, 4, 128, 256


def peaks():
    """(HBM GB/s, sustained bf16 TFLOP/s, source) from the driver-written MEASURED_PEAKS.json."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p["bf16_tflops_sustained"]), "measured (MEASURED_PEAKS.json)"
    except (OSError, KeyError, ValueError):
        return 6454.6, 1363.1, "B200 values measured earlier in round 1 (MEASURED_PEAKS.json absent)"


def timed(fn, iters, warmup, world):
    for _ in range(warmup):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item() * 1e-3


def config3(plan, world, tp_group, prompts, T, iters):
    dev = "cuda"
    h, hkv = plan.q_heads, plan.kv_heads
    Tt = prompts * T                                       # real tokens of this request shard
    gen = torch.Generator(device=dev).manual_seed(plan.dp_rank)
    qkv_g = torch.randn(Tt, (h + 2 * hkv) * D, generator=gen, device=dev).bfloat16()
    qkv_s = torch.randn(Tt, (h + 2 * hkv) * D, generator=gen, device=dev).bfloat16()
    cu = (torch.arange(prompts + 1, device=dev) * T).to(torch.int32)
    # GLA side: states of the prompts start from the shared meta state
    state = torch.zeros(prompts, hkv, D, D, device=dev, dtype=torch.bfloat16)
    meta_state = (torch.randn(hkv, D, D, generator=gen, device=dev) * 0.1).bfloat16()
    slots = torch.arange(prompts, device=dev, dtype=torch.int32)
    mode = torch.full((prompts,), 2, device=dev, dtype=torch.int32)
    # SWA side: pages, meta K/V, rotary table
    nb = Tt // BS
    kv = torch.zeros(2, nb, BS, hkv, D, device=dev, dtype=torch.bfloat16)
    slot_map = torch.randperm(nb, generator=gen, device=dev).repeat_interleave(BS) * BS + \
        torch.arange(BS, device=dev).repeat(nb)
    pos = torch.arange(T, device=dev).repeat(prompts)
    table = ops.rope_cos_sin_cache(D, T + M + 1, 1e6, device=dev)
    mk, mv = (torch.randn(M, hkv, D, generator=gen, device=dev).bfloat16() for _ in range(2))
    w1, w2 = torch.ones(h * D, device=dev), torch.ones(h * D, device=dev)
    o_gla = torch.empty(Tt, h, D, device=dev, dtype=torch.bfloat16)
    out = torch.empty(Tt, h * D, device=dev, dtype=torch.bfloat16)
    means = torch.empty(2, Tt, device=dev, dtype=torch.float32)
    split = lambda x: (x[:, :h * D].view(Tt, h, D), x[:, h * D:(h + hkv) * D].view(Tt, hkv, D),
                       x[:, (h + hkv) * D:].view(Tt, hkv, D))
    gq, gk, gv = split(qkv_g)
    sq, sk, sv = split(qkv_s)
    tp = plan.tp

    def layer():
        q2, k2, g2 = ops.gla_gate_70b(gq, gk)
        ops.gla_prefill_varlen(q2, k2, gv, g2, state, cu, slots, mode, meta_state=meta_state,
                               max_seqlen=T, out=o_gla)
        ops.rope_kv_write_paged(sq, sk, sv, table, pos, M, kv[0], kv[1], slot_map)
        o_swa = ops.swa_meta_prefill_varlen(sq, sk, sv, None, mk, mv, cu, cu, T, 4095)
        xa, xl = o_swa.view(Tt, h * D), o_gla.view(Tt, h * D)
        if tp > 1:     # one kernel: local means -> NVLink peer-memory exchange -> merge (f1)
            ops.branch_norm_merge(xa, xl, w1, w2, 1e-6, tp_world=tp, out=out, exchange=exchange)
        else:
            ops.branch_norm_merge(xa, xl, w1, w2, 1e-6, out=out)

    exchange = ops.BranchMergeExchange(tp_group, t_cap=Tt) if tp > 1 else None
    t = timed(layer, iters, 2, world)
    parts = {}
    parts["gla (gate+prefill)"] = timed(lambda: ops.gla_prefill_varlen(
        *ops.gla_gate_70b(gq, gk)[:2], gv, ops.gla_gate_70b(gq, gk)[2], state, cu, slots, mode,
        meta_state=meta_state, max_seqlen=T, out=o_gla), iters, 1, world)
    parts["swa (rope+kv+prefill)"] = timed(lambda: (
        ops.rope_kv_write_paged(sq, sk, sv, table, pos, M, kv[0], kv[1], slot_map),
        ops.swa_meta_prefill_varlen(sq, sk, sv, None, mk, mv, cu, cu, T, 4095)), iters, 1, world)
    n_keys = sum(M + min(i + 1, 4096) for i in range(T))
    swa_flops = prompts * h * 4 * D * n_keys
    return t, parts, plan.dp * Tt, swa_flops


def config5(plan, world, T, iters):
    dev = "cuda"
    n_req = len(plan.requests(8))
    hkv, h = plan.kv_heads, plan.q_heads
    if n_req == 0:
        return timed(lambda: None, iters, 1, world), 0
    gen = torch.Generator(device=dev).manual_seed(1)
    q = torch.randn(n_req * T, h, D, generator=gen, device=dev).bfloat16().relu()
    k = torch.randn(n_req * T, hkv, D, generator=gen, device=dev).bfloat16().relu()
    v = torch.randn(n_req * T, hkv, D, generator=gen, device=dev).bfloat16()
    g = (torch.nn.functional.logsigmoid(torch.randn(n_req * T, hkv, D, generator=gen, device=dev)) / 16).bfloat16()
    state = torch.zeros(n_req, hkv, D, D, device=dev, dtype=torch.bfloat16)
    cu = (torch.arange(n_req + 1, device=dev) * T).to(torch.int32)
    slots = torch.arange(n_req, device=dev, dtype=torch.int32)
    mode = torch.zeros(n_req, device=dev, dtype=torch.int32)
    out = torch.empty(n_req * T, h, D, device=dev, dtype=torch.bfloat16)
    t = timed(lambda: ops.gla_prefill_varlen(q, k, v, g, state, cu, slots, mode, max_seqlen=T, out=out),
              iters, 2, world)
    return t, n_req * T * (h * 2 * D * 2 + hkv * D * 6)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=3)
    ap.add_argument("--prompts", type=int, default=2)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    plan = sharding.plan(world, rank, H, HKV)
    tp_group = None
    if world > 1:
        dist.init_process_group("nccl")
        for d in range(plan.dp):                           # one group per head-shard set
            grp = dist.new_group([d * plan.tp + r for r in range(plan.tp)])
            if d == plan.dp_rank:
                tp_group = grp
    hbm, tfl, src = peaks()
    if a.config == 3:
        T = 32768
        t, parts, tokens, swa_flops = config3(plan, world, tp_group, a.prompts, T, a.iters)
        if rank == 0:
            print(json.dumps({"config": 3, "n_gpus": world, "plan": f"tp{plan.tp} x dp{plan.dp}",
                              "heads_per_gpu": [plan.q_heads, plan.kv_heads],
                              "prompts": plan.dp * a.prompts, "prompt_len": T, "meta_tokens": M,
                              "layer_ms": t * 1e3, "tok_s_per_layer": tokens / t,
                              "parts_ms": {k: v * 1e3 for k, v in parts.items()},
                              "swa_tflops_per_gpu": swa_flops / parts["swa (rope+kv+prefill)"] / 1e12,
                              "swa_frac_of_sustained_bf16": swa_flops / parts["swa (rope+kv+prefill)"] / 1e12 / tfl,
                              "exchange": "one all-reduce of 8*T bytes per layer" if plan.tp > 1 else "none",
                              "peaks": src}))
    else:
        rows = []
        for T in (2048, 4096, 8192, 16384, 32768, 65536, 131072):
            t, by = config5(plan, world, T, a.iters)
            rows.append({"T": T, "ms": round(t * 1e3, 3), "tok_s": 8 * T / t,
                         "gbs_per_gpu": by / t / 1e9, "hbm_frac_per_gpu": by / t / 1e9 / hbm})
        if rank == 0:
            print(json.dumps({"config": 5, "n_gpus": world, "plan": f"tp{plan.tp} x dp{plan.dp}",
                              "sequences": 8, "rows": rows, "peaks": src}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
