"""Extra measurement legs used by bench.py (``detail.vs_live`` and ``detail.configs``) and runnable
on their own (``python tools/bench_legs.py``).  Nothing here is on the product path.

* ``vs_live``       -- the GPU kernels the REFERENCE calls, as installed in this image and run on the
  same B200 in the same process: flash-attn 2.8.3 (``flash_attn_varlen_func``,
  ``flash_attn_with_kvcache``; FA2 mma.sync cubins for sm_100), flash-linear-attention 0.5.1
  (``chunk_gla``, Triton) and the reference's own Triton decode kernel
  (``my_fused_recurrent_fwd_kernel`` from ``baseline/_ref``), each timed next to the kernel of this
  repository that replaces it, at the BASELINE config shapes (SURVEY 2a "kernel to beat").
  Library calls happen ONLY in this leg.
* ``extra_configs`` -- one timed pass each of BASELINE configs 3 (70B hybrid layer, 32 K + meta,
  head-sharded, with f1's one all-reduce at TP > 1), 4 (1000 ragged decode requests) and 5 (GLA scan
  2 K..128 K).
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

D, H, HKV, BS = 128, 28, 4, 256


def _time(fn, iters=5, warmup=2):
    """Median CUDA-event time of ``fn`` in seconds."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3


def _time_graph(fn, reps=20):
    """Per-call time of a short kernel with launch overhead removed: ``reps`` calls in one graph."""
    fn()
    torch.cuda.synchronize()
    g, side = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps):
                fn()
    torch.cuda.synchronize()
    return _time(g.replay, iters=5, warmup=1) / reps


def _ragged_paged(lens, hkv, gen, dev):
    nblk = [(L + BS - 1) // BS for L in lens]
    nb = sum(nblk) + 1
    kc = torch.randn(nb, BS, hkv, D, generator=gen, device=dev).bfloat16()
    vc = torch.randn(nb, BS, hkv, D, generator=gen, device=dev).bfloat16()
    perm = torch.randperm(nb, generator=gen, device=dev).to(torch.int32)
    tables = torch.zeros(len(lens), max(nblk), dtype=torch.int32, device=dev)
    off = 0
    for i, n in enumerate(nblk):
        tables[i, :n] = perm[off:off + n]
        off += n
    return kc, vc, tables


def _row(ours_s, live_s, what, note=None):
    r = {"ours_us": ours_s * 1e6, "live_us": None if live_s is None else live_s * 1e6,
         "speedup": None if live_s is None else live_s / ours_s, "live": what}
    if note:
        r["note"] = note
    return r


def vs_live(dev="cuda", full=True):
    """dict: op -> {ours_us, live_us, speedup (live / ours, > 1 = this repository is faster)}."""
    from vllm_hymeta_b200 import ops
    res = {}
    gen = torch.Generator(device=dev).manual_seed(5)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
    W = 4096
    try:
        import flash_attn
    except Exception as e:                                   # pragma: no cover
        flash_attn = None
        res["flash_attn"] = {"unavailable": f"{type(e).__name__}: {e}"}

    # ---------------- SWA prefill, config 2: 64 x 8192, 28/4 heads, window 4096 + self
    N, L = (64, 8192) if full else (8, 8192)
    T = N * L
    q, k, v = (torch.randn(T, h, D, generator=gen, device=dev).bfloat16() for h in (H, HKV, HKV))
    cu = i32([i * L for i in range(N + 1)])
    out = torch.empty(T, H, D, device=dev, dtype=torch.bfloat16)
    t_our = _time(lambda: ops.swa_meta_prefill_varlen(q, k, v, None, None, None, cu, cu, L, W, out=out), 3, 1)
    t_live = None
    if flash_attn is not None:
        t_live = _time(lambda: flash_attn.flash_attn_varlen_func(q, k, v, cu, cu, L, L, softmax_scale=D ** -0.5,
                                                                 causal=True, window_size=(W, 0)), 3, 1)
    res[f"swa_prefill_{N}x{L}"] = _row(t_our, t_live, "flash_attn 2.8.3 flash_attn_varlen_func(window=(4096,0))")
    del out

    # ---------------- GLA prefill, config 2: fla chunk_gla on the repeat_kv'ed inputs (reference layout)
    qg, kg = torch.relu(q), torch.relu(k)
    g = (torch.nn.functional.logsigmoid(torch.randn(T, HKV, D, generator=gen, device=dev)) / 16).bfloat16()
    state = torch.zeros(N, HKV, D, D, dtype=torch.bfloat16, device=dev)
    slots = torch.arange(N, dtype=torch.int32, device=dev)
    mode = torch.zeros(N, dtype=torch.int32, device=dev)
    o = torch.empty(T, H, D, device=dev, dtype=torch.bfloat16)
    t_our = _time(lambda: ops.gla_prefill_varlen(qg, kg, v, g, state, cu, slots, mode, max_seqlen=L, out=o), 3, 1)
    del o
    t_live, note = None, None
    try:
        from fla.ops.gla import chunk_gla
        G = H // HKV
        rep = lambda x: x.view(N, L, HKV, D).repeat_interleave(G, dim=2)
        kr, vr, gr = rep(kg), rep(v), rep(g).float()
        q4 = qg.view(N, L, H, D)
        t_live = _time(lambda: chunk_gla(q4, kr, vr, gr, scale=D ** -0.5, output_final_state=True), 3, 1)
        note = "fla needs k/v/g repeated to 28 heads (as the reference's repeat_kv does); the repeat is not timed"
        del kr, vr, gr
    except Exception as e:
        note = f"fla chunk_gla not runnable: {type(e).__name__}: {str(e)[:200]}"
    res[f"gla_prefill_{N}x{L}"] = _row(t_our, t_live, "fla 0.5.1 chunk_gla (Triton)", note)
    del q, k, v, qg, kg, g
    torch.cuda.empty_cache()

    # ---------------- decode kernels at config 2 (B=64, ctx 8.2-8.7 K) and config 4 (B=1000 ragged)
    ref_mod = None
    try:
        from oracle import install_ref
        ref_mod = install_ref.load_ref_file("vllm_hymeta/model_for_7B/my_fused_recurrent.py", "_ref_mfr_bench")
    except Exception:
        ref_mod = None
    for tag, lens in (("B64_ctx8.2-8.7K", [8192 + 8 * i for i in range(64)]),
                      ("B1000_ragged_1K-9K", torch.randint(1024, 9217, (1000,), generator=torch.Generator().manual_seed(4)).tolist())):
        if not full and len(lens) > 64:
            continue
        B = len(lens)
        kc, vc, tables = _ragged_paged(lens, HKV, gen, dev)
        qd = torch.randn(B, H, D, generator=gen, device=dev).bfloat16()
        sl = i32(lens)
        od = torch.empty(B, H, D, device=dev, dtype=torch.bfloat16)
        t_our = _time_graph(lambda: ops.swa_meta_decode_paged(qd, kc, vc, tables, sl, None, None, W, out=od))
        t_live = None
        if flash_attn is not None:
            q4 = qd.unsqueeze(1)
            t_live = _time_graph(lambda: flash_attn.flash_attn_with_kvcache(
                q4, kc, vc, cache_seqlens=sl, block_table=tables, softmax_scale=D ** -0.5, causal=True,
                window_size=(W, 0)))
        res[f"swa_decode_{tag}"] = _row(t_our, t_live, "flash_attn 2.8.3 flash_attn_with_kvcache(block_table, window)")
        if flash_attn is not None and B == 64:
            # 70B form: + 128 meta keys.  reference = fwd_kvcache + non-causal meta pass + merge
            M = 128
            mk, mv = (torch.randn(M, HKV, D, generator=gen, device=dev).bfloat16() for _ in range(2))
            t_our = _time_graph(lambda: ops.swa_meta_decode_paged(qd, kc, vc, tables, sl, mk, mv, W - 1, out=od))
            mk4, mv4 = mk.unsqueeze(0), mv.unsqueeze(0)

            def two_pass():
                o1, l1 = flash_attn.flash_attn_with_kvcache(q4, kc, vc, cache_seqlens=sl, block_table=tables,
                                                            softmax_scale=D ** -0.5, causal=True,
                                                            window_size=(W - 1, 0), return_softmax_lse=True)
                o2, l2, _ = flash_attn.flash_attn_func(qd.unsqueeze(0), mk4, mv4, softmax_scale=D ** -0.5,
                                                       causal=False, return_attn_probs=True)
                o1, o2 = o1.float(), o2.transpose(0, 1).float()
                l1, l2 = l1, l2.permute(2, 1, 0)
                w = torch.sigmoid(l2 - l1).transpose(1, 2).unsqueeze(-1)
                return (o1 - w * (o1 - o2)).to(torch.bfloat16)
            try:
                t_live = _time_graph(two_pass)
                res["swa_meta_decode_B64_M128"] = _row(
                    t_our, t_live, "reference composition: flash_attn_with_kvcache + flash_attn_func (meta pass) + "
                                   "LSE merge in torch (fa_kernel_metatoken.py:124-194)")
            except Exception as e:
                res["swa_meta_decode_B64_M128"] = _row(t_our, None, "two-pass composition", f"{type(e).__name__}: {str(e)[:200]}")
        del kc, vc
        # GLA decode: ours compact (one state per kv-head) vs the reference Triton kernel in ITS layout
        gq = torch.randn(B, H, D, generator=gen, device=dev).bfloat16().relu()
        gk = torch.randn(B, HKV, D, generator=gen, device=dev).bfloat16().relu()
        gv = torch.randn(B, HKV, D, generator=gen, device=dev).bfloat16()
        gg = (torch.nn.functional.logsigmoid(torch.randn(B, HKV, D, generator=gen, device=dev)) / 16).bfloat16()
        st = torch.zeros(B, HKV, D, D, dtype=torch.bfloat16, device=dev)
        sidx = torch.arange(B, dtype=torch.int32, device=dev)
        og = torch.empty(B, H, D, device=dev, dtype=torch.bfloat16)
        t_our = _time_graph(lambda: ops.gla_decode(gq, gk, gv, gg, st, sidx, out=og))
        t_live, note = None, None
        if ref_mod is not None:
            try:
                G = H // HKV
                rep = lambda x: x.repeat_interleave(G, dim=1).unsqueeze(2).contiguous()
                q4r, k4r, v4r, g4r = gq.unsqueeze(2).contiguous(), rep(gk), rep(gv), rep(gg)
                st_ref = torch.zeros(B, H, D, D, dtype=torch.bfloat16, device=dev)
                t_live = _time_graph(lambda: ref_mod.my_fused_recurrent(q4r, k4r, v4r, g4r, st_ref, sidx))
                note = ("reference layout: 28 states per request (7x the state traffic of the compact layout), "
                        "repeat_kv of k/v/g not timed")
            except Exception as e:
                note = f"reference Triton kernel not runnable: {type(e).__name__}: {str(e)[:200]}"
        else:
            note = "baseline/_ref not present"
        res[f"gla_decode_{tag}"] = _row(t_our, t_live, "reference my_fused_recurrent_fwd_kernel (Triton)", note)
    return res


def extra_configs(dev="cuda", world=1, rank=0, tp_group=None, plan=None):
    """One timed pass each of BASELINE configs 3, 4, 5 on this rank's shard (device events)."""
    import json
    import torch.distributed as dist
    from vllm_hymeta_b200 import ops, sharding
    if plan is None:
        plan = sharding.plan(world, rank, H, HKV)
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        hbm, tfl = float(pk["hbm_gbs"]), float(pk["bf16_tflops_sustained"])
    except Exception:
        hbm, tfl = 6454.6, 1363.1
    out = {}
    gen = torch.Generator(device=dev).manual_seed(100 + plan.dp_rank)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
    h, hkv, tp = plan.q_heads, plan.kv_heads, plan.tp

    def tmax(t):
        if world > 1:
            x = torch.tensor([t], device=dev)
            dist.all_reduce(x, op=dist.ReduceOp.MAX)
            return x.item()
        return t

    # ---- config 3: one 70B hybrid layer, 2 prompts x 32 K + 128 meta tokens per request shard
    M, T, P = 128, 32768, 2
    Tt = P * T
    qkv_g = torch.randn(Tt, (h + 2 * hkv) * D, generator=gen, device=dev).bfloat16()
    qkv_s = torch.randn(Tt, (h + 2 * hkv) * D, generator=gen, device=dev).bfloat16()
    cu = i32([i * T for i in range(P + 1)])
    state = torch.zeros(P, hkv, D, D, device=dev, dtype=torch.bfloat16)
    meta_state = (torch.randn(hkv, D, D, generator=gen, device=dev) * 0.1).bfloat16()
    slots, mode = i32(list(range(P))), i32([2] * P)
    nb = Tt // BS
    kv = torch.zeros(2, nb, BS, hkv, D, device=dev, dtype=torch.bfloat16)
    slot_map = (torch.randperm(nb, generator=gen, device=dev).repeat_interleave(BS) * BS
                + torch.arange(BS, device=dev).repeat(nb))
    pos = torch.arange(T, device=dev).repeat(P)
    table = ops.rope_cos_sin_cache(D, T + M + 1, 1e6, device=dev)
    mk, mv = (torch.randn(M, hkv, D, generator=gen, device=dev).bfloat16() for _ in range(2))
    w1, w2 = torch.ones(h * D, device=dev), torch.ones(h * D, device=dev)
    o_gla = torch.empty(Tt, h, D, device=dev, dtype=torch.bfloat16)
    o_swa = torch.empty(Tt, h, D, device=dev, dtype=torch.bfloat16)
    merged = torch.empty(Tt, h * D, device=dev, dtype=torch.bfloat16)
    means = torch.empty(2, Tt, device=dev, dtype=torch.float32)
    sp = lambda x: (x[:, :h * D].view(Tt, h, D), x[:, h * D:(h + hkv) * D].view(Tt, hkv, D),
                    x[:, (h + hkv) * D:].view(Tt, hkv, D))
    gq, gk, gv = sp(qkv_g)
    sq, sk, sv = sp(qkv_s)

    def layer():
        q2, k2, g2 = ops.gla_gate_70b(gq, gk)
        ops.gla_prefill_varlen(q2, k2, gv, g2, state, cu, slots, mode, meta_state=meta_state, max_seqlen=T, out=o_gla,
                               sumsq_out=ss_gla)
        ops.rope_kv_write_paged(sq, sk, sv, table, pos, M, kv[0], kv[1], slot_map)
        ops.swa_meta_prefill_varlen(sq, sk, sv, None, mk, mv, cu, cu, T, 4095, out=o_swa, sumsq_out=ss_swa)
        xa, xl = o_swa.view(Tt, h * D), o_gla.view(Tt, h * D)
        if tp > 1 and exchange is not None:   # one kernel: local means -> NVLink peer-memory exchange -> merge (f1)
            ops.branch_norm_merge(xa, xl, w1, w2, 1e-6, tp_world=tp, out=merged, exchange=exchange,
                                  sumsq=(ss_swa, ss_gla))
        elif tp > 1:
            ops.branch_norm_merge(xa, xl, w1, w2, 1e-6, tp_world=tp, out=merged, group=tp_group)
        else:
            ops.branch_norm_merge(xa, xl, w1, w2, 1e-6, out=merged)
    # f1 as specified: the attention epilogues emit the per-(token, head) sums of squares the merge needs
    ss_swa = torch.empty(Tt, h, dtype=torch.float32, device=dev) if tp > 1 else None
    ss_gla = torch.empty(Tt, h, dtype=torch.float32, device=dev) if tp > 1 else None
    exchange = None
    if tp > 1:
        # the peer-memory exchange is collective: every rank of the job agrees on whether it came up
        # (symmetric memory needs P2P access inside the TP group); otherwise all fall back to the NCCL form
        try:
            exchange = ops.BranchMergeExchange(tp_group, t_cap=Tt)
            ok = 1
        except Exception:
            ok = 0
        flag = torch.tensor([ok], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if flag.item() == 0:
            exchange = None
    if world > 1:
        dist.barrier()
    t = tmax(_time(layer, 3, 1))
    n_keys = sum(M + min(i + 1, 4096) for i in range(T))
    out["config3_70b_hybrid_layer_32k"] = {
        "plan": f"tp{plan.tp} x dp{plan.dp}", "heads_per_gpu": [h, hkv], "prompts": plan.dp * P, "prompt_len": T,
        "meta_tokens": M, "layer_ms": t * 1e3, "tok_s_per_layer": plan.dp * Tt / t,
        "exchange": ("none" if tp == 1 else "fused branch-merge kernel: sum-of-squares partials from the attention epilogues, 16 B per token and peer over NVLink peer memory (f1)"
                     if exchange is not None else "one NCCL all-reduce of 8*T bytes per layer (f1; peer memory unavailable)"),
        "swa_flops_per_gpu": P * h * 4 * D * n_keys}
    del qkv_g, qkv_s, kv, o_gla, o_swa, merged, slot_map
    torch.cuda.empty_cache()

    # ---- config 4: 1000 ragged decode requests (this rank's request / head shard), 7B layer pair
    all_lens = torch.randint(1024, 9217, (1000,), generator=torch.Generator().manual_seed(4)).tolist()
    lens = [all_lens[i] for i in plan.requests(1000)]
    B = len(lens)
    kc, vc, tables = _ragged_paged(lens, hkv, gen, dev)
    qd = torch.randn(B, h, D, generator=gen, device=dev).bfloat16()
    sl = i32(lens)
    od = torch.empty(B, h, D, device=dev, dtype=torch.bfloat16)
    t_swa = tmax(_time_graph(lambda: ops.swa_meta_decode_paged(qd, kc, vc, tables, sl, None, None, 4096, out=od), 10))
    gk_ = torch.randn(B, hkv, D, generator=gen, device=dev).bfloat16().relu()
    gv_ = torch.randn(B, hkv, D, generator=gen, device=dev).bfloat16()
    gg_ = (torch.nn.functional.logsigmoid(torch.randn(B, hkv, D, generator=gen, device=dev)) / 16).bfloat16()
    st = torch.zeros(B, hkv, D, D, dtype=torch.bfloat16, device=dev)
    sidx = torch.arange(B, dtype=torch.int32, device=dev)
    t_gla = tmax(_time_graph(lambda: ops.gla_decode(qd.relu(), gk_, gv_, gg_, st, sidx, out=od), 10))
    swa_bytes = sum(hkv * 2 * D * 2 * min(L, 4097) for L in lens) + B * (h * 2 * D * 2 + hkv * 2 * D * 2)
    gla_bytes = B * (hkv * 2 * D * D * 2 + h * 2 * D * 2 + hkv * D * 6)
    out["config4_decode_1000_ragged"] = {
        "requests_this_gpu": B, "heads_per_gpu": [h, hkv],
        "swa_decode_us": t_swa * 1e6, "swa_gbs": swa_bytes / t_swa / 1e9, "swa_hbm_frac": swa_bytes / t_swa / 1e9 / hbm,
        "gla_decode_us": t_gla * 1e6, "gla_gbs": gla_bytes / t_gla / 1e9, "gla_hbm_frac": gla_bytes / t_gla / 1e9 / hbm,
        "note": "gla_decode_us includes a relu of q in torch (inputs are otherwise pre-gated)"}
    del kc, vc
    torch.cuda.empty_cache()

    # ---- config 5: GLA scan, 8 sequences x T (strong scaling: chains spread over the GPUs)
    rows = []
    n_req = len(plan.requests(8))
    for T5 in (2048, 8192, 32768, 131072):
        if n_req == 0:
            rows.append({"T": T5, "ms": tmax(0.0) * 1e3})
            continue
        n = n_req * T5
        q5 = torch.randn(n, h, D, generator=gen, device=dev).bfloat16().relu()
        k5 = torch.randn(n, hkv, D, generator=gen, device=dev).bfloat16().relu()
        v5 = torch.randn(n, hkv, D, generator=gen, device=dev).bfloat16()
        g5 = (torch.nn.functional.logsigmoid(torch.randn(n, hkv, D, generator=gen, device=dev)) / 16).bfloat16()
        st5 = torch.zeros(n_req, hkv, D, D, device=dev, dtype=torch.bfloat16)
        o5 = torch.empty(n, h, D, device=dev, dtype=torch.bfloat16)
        cu5 = i32([i * T5 for i in range(n_req + 1)])
        t5 = tmax(_time(lambda: ops.gla_prefill_varlen(q5, k5, v5, g5, st5, cu5, i32(list(range(n_req))),
                                                       i32([0] * n_req), max_seqlen=T5, out=o5), 3, 1))
        by = n * (h * 2 * D * 2 + hkv * D * 6)
        rows.append({"T": T5, "ms": t5 * 1e3, "tok_s": 8 * T5 / t5, "hbm_frac_per_gpu": by / t5 / 1e9 / hbm})
        del q5, k5, v5, g5, o5
    out["config5_gla_scan_8seq"] = {"plan": f"tp{plan.tp} x dp{plan.dp}", "rows": rows}
    out["peaks"] = {"hbm_gbs": hbm, "bf16_tflops_sustained": tfl}
    c3 = out["config3_70b_hybrid_layer_32k"]
    c3["note"] = "layer = gate + GLA prefill (meta-state init) + fused RoPE/KV write + SWA-meta prefill + branch merge"
    return out


if __name__ == "__main__":
    import json
    torch.cuda.set_device(0)
    print(json.dumps({"vs_live": vs_live(full="--small" not in sys.argv)}, indent=1))
    print(json.dumps({"configs": extra_configs()}, indent=1))
