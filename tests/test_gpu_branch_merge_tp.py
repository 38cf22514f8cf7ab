"""f1 at tensor-parallel world size 2 on real GPUs (skips on a single-GPU box): the fused
compute -> NVLink exchange -> merge kernel (``hymeta_branch_norm_merge_tp`` through
``ops.branch_norm_merge(exchange=...)``) and the NCCL form (``branch_sumsq`` + all-reduce + merge)
against the oracle's HymetaRMSNormTP restatement on every rank's hidden slice."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank: int, world: int, port: int, q):
    try:
        sys.path.insert(0, ROOT)
        import torch.distributed as dist
        from oracle import gla as ogla                      # checker only
        from vllm_hymeta_b200 import ops
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                                device_id=torch.device("cuda", rank))
        eps = 1e-6
        ex = ops.BranchMergeExchange(None, t_cap=4096)                      # epoch in device memory
        ex_host = ops.BranchMergeExchange(None, t_cap=4096, device_epoch=False)   # epoch by value
        worst = 0.0
        for T, hidden in ((1, 3584), (257, 3584), (4096, 3584), (64, 2048)):
            gen = torch.Generator().manual_seed(T)           # same data on every rank
            xa = (torch.randn(T, hidden, generator=gen) * 3).bfloat16()
            xl = (torch.randn(T, hidden, generator=gen) * 0.2).bfloat16()
            w1 = 1 + 0.1 * torch.randn(hidden, generator=gen)
            w2 = 1 + 0.1 * torch.randn(hidden, generator=gen)
            hl = hidden // world
            sl = [slice(r * hl, (r + 1) * hl) for r in range(world)]
            s = sl[rank]
            tot = [sum(x[:, qq].float().pow(2).mean(-1, keepdim=True, dtype=torch.float32) for qq in sl)
                   for x in (xa, xl)]
            ref = ogla.branch_norm_merge(xa[:, s], xl[:, s], w1[s], w2[s], eps, world,
                                         (lambda v: tot[0], lambda v: tot[1]))
            dxa, dxl = xa[:, s].cuda().contiguous(), xl[:, s].cuda().contiguous()
            dw1, dw2 = w1[s].cuda().contiguous(), w2[s].cuda().contiguous()
            for rep in range(3):                             # epochs advance, both slot parities get used
                fused = ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world, exchange=ex)
            nccl = ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world)
            for rep in range(2):
                fused_host = ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world, exchange=ex_host)
            assert torch.equal(fused_host, fused), (T, hidden)
            # f1 as specified: the per-(token, head) partials the attention epilogues emit feed the merge
            heads = hl // 128
            sa = dxa.float().view(T, heads, 128).pow(2).sum(-1).contiguous()
            sb = dxl.float().view(T, heads, 128).pow(2).sum(-1).contiguous()
            fused_ss = ops.branch_norm_merge(dxa, dxl, dw1, dw2, eps, tp_world=world, exchange=ex, sumsq=(sa, sb))
            torch.cuda.synchronize()
            d_ss = (fused_ss.float() - fused.float()).abs()
            assert (d_ss > 0).float().mean().item() < 0.01 and \
                (d_ss <= 2.0 ** -6 * fused.float().abs().clamp(min=1.0)).all(), (T, hidden, d_ss.max().item())
            for name, got in (("fused", fused), ("nccl", nccl)):
                diff = (got.float().cpu() - ref.float()).abs()
                bad = int((diff > 2.0 ** -6 * ref.float().abs().clamp(min=1.0)).sum())
                assert bad == 0, (name, T, hidden, bad, diff.max().item())
                worst = max(worst, diff.max().item())
            # world 2: a + b in either order is the same float, so the two forms agree bit for bit
            if world == 2:
                assert torch.equal(fused, nccl), (T, hidden)
        # ---- CUDA graph: the captured launch is replayed with new inputs, the kernel advances its
        #      own epoch; eager calls in between keep the same epoch sequence on both ranks
        T, hl = 64, 3584 // world
        gen = torch.Generator().manual_seed(99)
        xs = [(torch.randn(T, 3584, generator=gen) * (1 + i)).bfloat16()[:, rank * hl:(rank + 1) * hl].cuda().contiguous()
              for i in range(4)]
        wa = (1 + 0.1 * torch.randn(3584, generator=gen))[rank * hl:(rank + 1) * hl].cuda().contiguous()
        wb = (1 + 0.1 * torch.randn(3584, generator=gen))[rank * hl:(rank + 1) * hl].cuda().contiguous()
        sa, sb = torch.empty_like(xs[0]), torch.empty_like(xs[0])
        gout = torch.empty_like(xs[0])
        sa.copy_(xs[0]); sb.copy_(xs[1])
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            ops.branch_norm_merge(sa, sb, wa, wb, eps, tp_world=world, exchange=ex, out=gout)   # warm-up
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            ops.branch_norm_merge(sa, sb, wa, wb, eps, tp_world=world, exchange=ex, out=gout)
            ops.branch_norm_merge(gout, sb, wa, wb, eps, tp_world=world, exchange=ex, out=gout)  # two calls per replay
        for i in range(5):
            a, b = xs[i % 4], xs[(i + 1) % 4]
            sa.copy_(a); sb.copy_(b)
            graph.replay()
            want = ops.branch_norm_merge(a, b, wa, wb, eps, tp_world=world)          # NCCL form, eager
            want = ops.branch_norm_merge(want, b, wa, wb, eps, tp_world=world)
            if i % 2:
                ops.branch_norm_merge(a, b, wa, wb, eps, tp_world=world, exchange=ex)  # eager call between replays
            torch.cuda.synchronize()
            assert torch.equal(gout, want), ("graph replay", i)
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok", worst))
    except Exception as e:  # pragma: no cover - reported to the parent
        import traceback
        q.put((rank, "fail", traceback.format_exc()))


def test_branch_merge_tp2_fused_and_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 1000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = []
    for _ in procs:
        res.append(q.get(timeout=300))
    for p in procs:
        p.join(timeout=60)
    for rank, status, info in res:
        assert status == "ok", f"rank {rank}:\n{info}"
