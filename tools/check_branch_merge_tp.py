#!/usr/bin/env python3
"""N-GPU check of the one exchange adjacent to the path: the 70B branch merge with
the hidden size sharded over the tensor-parallel ranks, per-token means summed by ONE NCCL
all-reduce of 8*T bytes.  Every rank checks its slice against the oracle restatement of
HymetaRMSNormTP evaluated on the full (replicated, same seed) activations.
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/check_branch_merge_tp.py"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from oracle import gla as ogla                      # checker only
from vllm_hymeta_b200 import ops


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("nccl")
    T, hidden, eps = int(os.environ.get("BM_T", "65536")), 3584, 1e-6
    gen = torch.Generator().manual_seed(0)           # same data on every rank
    xa = (torch.randn(T, hidden, generator=gen) * 3).bfloat16()
    xl = (torch.randn(T, hidden, generator=gen) * 0.2).bfloat16()
    w1 = 1 + 0.1 * torch.randn(hidden, generator=gen)
    w2 = 1 + 0.1 * torch.randn(hidden, generator=gen)
    hl = hidden // world
    s = slice(rank * hl, (rank + 1) * hl)
    sl = [slice(r * hl, (r + 1) * hl) for r in range(world)]
    tot = [sum(x[:, q].float().pow(2).mean(-1, keepdim=True, dtype=torch.float32) for q in sl) for x in (xa, xl)]
    ref = ogla.branch_norm_merge(xa[:, s], xl[:, s], w1[s], w2[s], eps, world,
                                 (lambda v: tot[0], lambda v: tot[1]))
    dxa, dxl = xa[:, s].cuda().contiguous(), xl[:, s].cuda().contiguous()
    dw1, dw2 = w1[s].cuda().contiguous(), w2[s].cuda().contiguous()
    got = ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world)      # all-reduce inside
    ex = ops.BranchMergeExchange(None, t_cap=T) if world > 1 else None
    if ex is not None:
        fused = ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world, exchange=ex)
        torch.cuda.synchronize()
        dfu = (fused.float() - got.float()).abs().max().item()
        if rank == 0:
            print(f"fused NVLink-exchange kernel vs NCCL form: max|d| = {dfu:.3e}", flush=True)
    torch.cuda.synchronize()
    diff = (got.float().cpu() - ref.float()).abs()
    bad = int((diff > 2.0 ** -6 * ref.float().abs().clamp(min=1.0)).sum())
    neq = float((diff > 0).float().mean())
    # timing: sumsq + all-reduce + merge, device events, max over ranks
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(20):
        ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world)
    ev1.record()
    torch.cuda.synchronize()
    us = torch.tensor([ev0.elapsed_time(ev1) / 20 * 1e3], device="cuda")
    if world > 1:
        dist.all_reduce(us, op=dist.ReduceOp.MAX)
    print(f"rank {rank}/{world}: slice [{s.start}:{s.stop}] bad={bad} differing={neq:.3%} "
          f"sumsq + NCCL all-reduce + merge {us.item():.1f} us (max over ranks, T={T})", flush=True)
    if ex is not None:
        for Tn in (64, 4096, T):
            for name, fn in (("NCCL form ", lambda: ops.branch_norm_merge(dxa[:Tn], dxl[:Tn], dw1, dw2, eps, tp_world=world)),
                             ("fused form", lambda: ops.branch_norm_merge(dxa[:Tn], dxl[:Tn], dw1, dw2, eps, tp_world=world,
                                                                          exchange=ex))):
                for _ in range(3):
                    fn()
                dist.barrier()
                torch.cuda.synchronize()
                ev0.record()
                for _ in range(20):
                    fn()
                ev1.record()
                torch.cuda.synchronize()
                u = torch.tensor([ev0.elapsed_time(ev1) / 20 * 1e3], device="cuda")
                dist.all_reduce(u, op=dist.ReduceOp.MAX)
                if rank == 0:
                    by = Tn * hl * 2 * 3
                    print(f"  T={Tn:6d} {name}: {u.item():8.1f} us  ({by / u.item() / 1e3:7.1f} GB/s of 2R+1W per GPU)", flush=True)
    graph_ok = True
    if ex is not None:
        # CUDA-graph replay of the fused form (device-side epoch): the decode-step use, T = 64
        Tn = 64
        sa, sb, gout = dxa[:Tn].clone(), dxl[:Tn].clone(), torch.empty_like(dxa[:Tn])
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            ops.branch_norm_merge(sa, sb, dw1, dw2, eps, tp_world=world, exchange=ex, out=gout)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            ops.branch_norm_merge(sa, sb, dw1, dw2, eps, tp_world=world, exchange=ex, out=gout)
        for i in range(4):
            sa.copy_(dxa[i * Tn:(i + 1) * Tn]); sb.copy_(dxl[i * Tn:(i + 1) * Tn])
            graph.replay()
            want = ops.branch_norm_merge(sa, sb, dw1, dw2, eps, tp_world=world)
            torch.cuda.synchronize()
            d = (gout.float() - want.float()).abs()
            graph_ok &= bool((d <= 2.0 ** -6 * want.float().abs().clamp(min=1.0)).all())
        dist.barrier()
        torch.cuda.synchronize()
        ev0.record()
        for _ in range(50):
            graph.replay()
        ev1.record()
        torch.cuda.synchronize()
        u = torch.tensor([ev0.elapsed_time(ev1) / 50 * 1e3], device="cuda")
        dist.all_reduce(u, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(f"  T={Tn:6d} fused form, CUDA-graph replays: {u.item():8.1f} us  (results match the NCCL form: {graph_ok})", flush=True)
    ok = torch.tensor([int(bad == 0 and neq <= 0.02 and graph_ok)], device="cuda")
    if world > 1:
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        dist.destroy_process_group()
    sys.exit(0 if ok.item() else 1)


if __name__ == "__main__":
    main()
