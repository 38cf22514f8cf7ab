#!/usr/bin/env python
"""Contract benchmark: the Hymeta-7B attention stack (14 GLA + 14 SWA layers), batch 64,
prefill 8192 tokens + 512 decode steps per request (BASELINE.json configs[1]) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one pass of the whole hot path over one synthetic batch, driven through the
REFERENCE-FACING plugin classes (the calls vLLM / the reference model code make):

  prefill  14 x MetaAttentionImpl.forward            (paged-KV write + SWA prefill, post-RoPE q/k/v)
           14 x GatedLinearAttention.forward         (gate pre-processing + GLA chunked prefill with
                                                      slot states + g_norm epilogue, raw qkv_proj rows)
  decode   512 x [ 14 x MetaAttentionImpl.forward    (ONE launch: KV write + split-K paged decode)
                   14 x GatedLinearAttention.forward (gate + recurrent state update + g_norm)
                   MetaAttentionMetadata-style advance_step on device ]   -- replayed from CUDA graphs

`value` times that step with the inputs resident in HBM.  `e2e` times the SAME step from HOST
buffers: per step the activations are copied from pinned memory, the slot manager admits the 64
requests and releases the previous ones, MetaAttentionMetadataBuilder builds the prefill and decode
metadata on the host and ships it, and the last decode outputs are read back.

N > 1 (torchrun, one rank per GPU): heads are sharded min(N,4)-way exactly as vLLM TP shards them
(28/4 q/kv heads -> 14/2 -> 7/1) and requests are sharded beyond that (N=8: 4 head shards x 2
request shards); the global batch is 64*N so per-GPU work is constant (weak scaling).  The path has
no exchange step, hence no collective in the data path.  `detail.configs` adds one timed pass of
BASELINE configs 3 (with its one f1 all-reduce at TP > 1), 4 and 5; `detail.vs_live` (N = 1) times the
library kernels the reference calls next to ours on the same GPU.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H_FULL, HKV_FULL, D = 28, 4, 128
N_GLA_LAYERS = N_SWA_LAYERS = 14
WINDOW_LEFT = 4096          # 7B: window_size=(W, 0) -> W keys before the query
BLOCK_SIZE = 256
METRIC = "attention-stack throughput (prefill + decode tokens)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--prefill", type=int, default=8192)
    ap.add_argument("--decode", type=int, default=512)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip detail.vs_live / detail.configs")
    return ap.parse_args()


def workload_config(args, n):
    tp = min(n, 4)
    return {"workload": f"hymeta7b_attention_stack_{N_GLA_LAYERS}gla+{N_SWA_LAYERS}swa_"
                        f"prefill{args.prefill}_decode{args.decode}_batch{args.batch}_per_gpu",
            "global_batch": args.batch * n, "prefill_len": args.prefill, "decode_steps": args.decode,
            "heads_q_kv": [H_FULL, HKV_FULL], "head_dim": D, "window_left": WINDOW_LEFT,
            "kv_block_size": BLOCK_SIZE, "parallelism": f"tp{tp}(heads) x dp{n // tp}(requests)",
            "l2": "inputs larger than L2 (>= 5 GB streamed per layer; decode KV 0.5 GB per layer-step)"}


# ----------------------------------------------------------------------------------------------
# CPU arm: the reference's torch path on the host cores (oracle port of the hot-path arithmetic)
# ----------------------------------------------------------------------------------------------
CPU_SAMPLE = (8192, 16)     # prefill tokens (>= the 4096-key window, = the workload's prompt), decode steps
CPU_DECODE_BATCH = 16


def cpu_sample_desc(pt, ds):
    return (f"oracle port of the reference torch path (fp32, all host threads): ONE GLA layer + ONE SWA "
            f"layer at the 7B shape (28/4 heads, window {WINDOW_LEFT} + self): prefill of one {pt}-token "
            f"prompt (step recurrence; blocked dense softmax, the window binds) + {ds} decode steps of a "
            f"batch of {CPU_DECODE_BATCH} requests at context > {pt} (batched tensor ops); per-token prefill "
            f"and decode costs are combined at the workload's prefill:decode token mix and divided by 14 "
            f"layer pairs")


def cpu_baseline(args, threads: int, reps: int = 1):
    from oracle import ref_cpu
    pt, ds = CPU_SAMPLE
    best = None
    for _ in range(reps):
        s = ref_cpu.port_stack_sample(pt, ds, threads, decode_batch=CPU_DECODE_BATCH)
        if best is None or s["total_s"] < best["total_s"]:
            best = s
    out = {"value": ref_cpu.stack_tok_s(best, args.prefill, args.decode, N_GLA_LAYERS), "unit": "tok/s",
           "cores": threads, "kind": "port", "sample": cpu_sample_desc(pt, ds),
           "sample_seconds": {k: round(v, 4) for k, v in best.items() if k.endswith("_s")}}
    try:    # BASELINE.json configs[0] through the reference's own hf_7B_model modules (baseline/_ref)
        out["config1_reference_modules"] = ref_cpu.reference_config1(threads)
    except Exception as e:                                   # never fatal for the bench line
        out["config1_reference_modules"] = {"unavailable": f"{type(e).__name__}: {e}"}
    return out


def run_reference(args):
    """`--impl reference`: the reference's CPU implementation of the path on the host cores.  Each
    step is ONE bounded sample of the workload (see cpu_sample_desc); value = the workload's tok/s
    implied by the mean step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref_cpu
    threads = os.cpu_count() or 1
    pt, ds = CPU_SAMPLE
    for _ in range(args.warmup):
        ref_cpu.port_stack_sample(512, 2, threads, decode_batch=2)           # warm-up: allocator, thread pool
    acc, t0 = None, time.perf_counter()
    for _ in range(args.steps):
        s = ref_cpu.port_stack_sample(pt, ds, threads, decode_batch=CPU_DECODE_BATCH)
        acc = s if acc is None else {k: acc[k] + v if k.endswith("_s") else v for k, v in s.items()}
    dt = (time.perf_counter() - t0) / args.steps
    mean = {k: v / args.steps if k.endswith("_s") else v for k, v in acc.items()}
    value = ref_cpu.stack_tok_s(mean, args.prefill, args.decode, N_GLA_LAYERS)
    cfg = workload_config(args, 1)          # the workload whose throughput the sample estimates
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "tok/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": "tok/s", "cores": threads, "kind": "port",
                         "sample": "each step: " + cpu_sample_desc(pt, ds),
                         "sample_seconds": {k: round(v, 4) for k, v in mean.items() if k.endswith("_s")}},
        "e2e": {"value": value, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "sample": cpu_sample_desc(pt, ds),
        "note": "ms_per_step is the wall time of one bounded sample, not of the full workload; value is the "
                "workload's tok/s implied by the sample's per-token prefill and decode costs",
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for t, line in self.rows:
            if t < t0 or t > t1:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                  "sw_power_cap"), f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_peaks():
    peaks = {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0, "bf16_tflops": 1650.0,
             "src": "fallback (B200_PROFILING.md)"}
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peaks = {"hbm_gbs": pk["hbm_gbs"], "bf16_tflops_sustained": pk["bf16_tflops_sustained"],
                 "bf16_tflops": pk.get("bf16_tflops", pk["bf16_tflops_sustained"]),
                 "src": "measured (MEASURED_PEAKS.json)"}
    except Exception:
        pass
    return peaks


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from types import SimpleNamespace
    from vllm_hymeta_b200 import ops, sharding
    from vllm_hymeta_b200.attention.meta_attention import (MetaAttentionImpl,
                                                           MetaAttentionMetadataBuilder, SeqData)
    from vllm_hymeta_b200.model_for_7B.gla_attention import GatedLinearAttention
    from vllm_hymeta_b200.model_for_7B.gla_cache import GLACacheManager, GLACacheParams

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = max(world, 1)
    plan = sharding.plan(n, rank, H_FULL, HKV_FULL)
    tp = plan.tp
    H, Hkv = plan.q_heads, plan.kv_heads
    B = args.batch * tp            # requests resident on this GPU (its request shard)
    Tp, Sd = args.prefill, args.decode
    T = B * Tp
    bs = BLOCK_SIZE
    bf = torch.bfloat16
    torch.manual_seed(1234 + rank)

    # ---- synthetic inputs, resident in HBM: what the model code hands to the plugin
    def randn(*shape):
        return torch.randn(*shape, device=dev, dtype=bf)
    QKV = (H + 2 * Hkv) * D
    swa_q, swa_k, swa_v = randn(T, H, D), randn(T, Hkv, D), randn(T, Hkv, D)      # post-RoPE q/k/v
    gla_qkv, gla_gk = randn(T, QKV), randn(T, Hkv * D)                             # raw qkv_proj / gk_proj rows
    out_swa = torch.empty(T, H, D, device=dev, dtype=bf)
    dq, dk, dv = randn(B, H, D), randn(B, Hkv, D), randn(B, Hkv, D)
    dg_qkv, dg_gk = randn(B, QKV), randn(B, Hkv * D)
    dout_swa = torch.empty(B, H, D, device=dev, dtype=bf)

    # ---- the plugin objects: backend impl + 7B GLA core + slot-state manager + metadata builder
    impl = MetaAttentionImpl(H, D, D ** -0.5, Hkv, None, WINDOW_LEFT + 1, "auto", num_meta_tokens=0)
    layers = [SimpleNamespace(kv_already_written=False) for _ in range(N_SWA_LAYERS)]
    gla_core = GatedLinearAttention(H_FULL * D, H_FULL, HKV_FULL, tp_size=tp, tp_rank=plan.tp_rank,
                                    g_norm_weight=torch.ones(D, device=dev, dtype=bf))
    mgr = GLACacheManager(bf, (N_GLA_LAYERS, B, Hkv, D, D), device=dev)
    gla_state = mgr.cache
    blocks_per_seq = (Tp + Sd + bs - 1) // bs
    nb = B * blocks_per_seq
    kv_caches = [torch.zeros(2, nb, bs, Hkv, D, device=dev, dtype=bf) for _ in range(N_SWA_LAYERS)]
    gen = torch.Generator(device="cpu").manual_seed(7)
    tables_h = torch.randperm(nb, generator=gen).to(torch.int32).view(B, blocks_per_seq)
    table_rows = tables_h.tolist()

    def build_prefill_md():
        b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device=dev, pin_memory=True)
        for row in table_rows:
            b.add(SeqData(True, Tp, Tp, 0, row))
        return b.build()

    def build_decode_md():
        b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device=dev, pin_memory=True)
        for row in table_rows:
            b.add(SeqData(False, Tp + 1, 1, Tp, row))
        return b.build()

    step_no = [0]

    def admit_requests():
        """Slot manager: release the previous batch, admit B new requests -> slot indices [B]."""
        k = step_no[0]
        step_no[0] += 1
        ids = {f"req-{k}-{i}": [i] for i in range(B)}
        done = [f"req-{k - 1}-{i}" for i in range(B)] if k > 0 else []
        _, idx = mgr.current_run_tensors(request_ids_to_seq_ids=ids, finished_requests_ids=done)
        return idx

    md_p = build_prefill_md()
    md_d = build_decode_md()                       # static buffers the decode graphs read
    idx_buf = admit_requests().clone()             # static slot-index buffer
    seq0, slot0 = md_d.seq_lens_tensor.clone(), md_d.slot_mapping.clone()

    def decode_swa():                              # one fused launch per layer-step
        for l in range(N_SWA_LAYERS):
            impl.forward(layers[l], dq, dk, dv, None, None, None, kv_caches[l], md_d, output=dout_swa)

    def decode_gla():
        o = None
        for l in range(N_GLA_LAYERS):
            o = gla_core.forward(dg_qkv, dg_gk, GLACacheParams(gla_state[l], idx_buf), md_d)
        # multi-step scheduling: the metadata of the next step is derived on the device
        ops.advance_step(B, B, bs, md_d.seq_lens_tensor, md_d.slot_mapping, md_d.block_tables)
        return o

    # ---- CUDA graphs for the launch-bound decode step (SWA half, GLA half + advance)
    decode_swa(); decode_gla()                     # eager once: sizes workspaces
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    l0 = ops.launch_count()
    g_swa, g_gla = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g_swa, stream=side):
            decode_swa()
        l1 = ops.launch_count()
        with torch.cuda.graph(g_gla, stream=side):
            dout_gla = decode_gla()
    l2 = ops.launch_count()
    launches_swa_step, launches_gla_step = l1 - l0, l2 - l1
    torch.cuda.synchronize()

    ev = lambda: torch.cuda.Event(enable_timing=True)

    def one_step(rec=None, md=None, idx=None, before_gla=None, fresh_md_d=None, swa_in=None):
        sq, sk, sv = (swa_q, swa_k, swa_v) if swa_in is None else swa_in
        md = md_p if md is None else md
        idx = idx_buf if idx is None else idx
        if fresh_md_d is None:                     # same request set as before: rewind the decode metadata
            md_d.seq_lens_tensor.copy_(seq0)
            md_d.slot_mapping.copy_(slot0)
        else:                                      # host-built metadata -> the graphs' static buffers
            md_d.seq_lens_tensor.copy_(fresh_md_d.seq_lens_tensor)
            md_d.slot_mapping.copy_(fresh_md_d.slot_mapping)
            md_d.block_tables.copy_(fresh_md_d.block_tables)
            idx_buf.copy_(idx)
        e = [ev() for _ in range(4)] if rec is not None else None
        if e: e[0].record()
        for l in range(N_SWA_LAYERS):
            impl.forward(layers[l], sq, sk, sv, None, None, None, kv_caches[l], md, output=out_swa)
        if e: e[1].record()
        if before_gla is not None:
            before_gla()
        for l in range(N_GLA_LAYERS):
            gla_core.forward(gla_qkv, gla_gk, GLACacheParams(gla_state[l], idx), md)
        if e: e[2].record()
        if rec is None:
            for _ in range(Sd):
                g_swa.replay(); g_gla.replay()
        else:
            marks = []
            for _ in range(Sd):
                a, b, c = ev(), ev(), ev()
                a.record(); g_swa.replay(); b.record(); g_gla.replay(); c.record()
                marks.append((a, b, c))
            e[3].record()
            rec.append((e, marks))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_step()
    barrier()
    launches0 = ops.launch_count()
    sampler = ClockSampler(local) if rank == 0 else None
    rec = []
    ops.timing_begin()                             # per-launch events of the eager (prefill) phase
    t_wall0 = time.perf_counter()
    start, stop = ev(), ev()
    start.record()
    for _ in range(args.steps):
        one_step(rec)
    stop.record()
    barrier()
    t_wall1 = time.perf_counter()
    launch_events = ops.timing_end()
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    eager_launches = ops.launch_count() - launches0
    elapsed = start.elapsed_time(stop) * 1e-3
    if world > 1:
        t = torch.tensor([elapsed], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed = t.item()
    step_s = elapsed / args.steps
    tokens_per_gpu = B * (Tp + Sd) / tp       # a head shard does 1/tp of each token's work
    total_tokens = args.batch * n * (Tp + Sd)
    value = total_tokens / step_s
    gpu_launches = eager_launches + args.steps * Sd * (launches_swa_step + launches_gla_step)

    # ---- phase timings (events recorded inside the timed region) and per-launch kernel times
    ms = lambda a, b: a.elapsed_time(b)
    t_pswa = sum(ms(e[0], e[1]) for e, _ in rec) / args.steps
    t_pgla = sum(ms(e[1], e[2]) for e, _ in rec) / args.steps
    t_dec = sum(ms(e[2], e[3]) for e, _ in rec) / args.steps
    t_dswa = sum(ms(a, b) for _, m in rec for a, b, c in m) / args.steps
    t_dgla = sum(ms(b, c) for _, m in rec for a, b, c in m) / args.steps
    per_launch = {}
    for name, a, b in launch_events:
        per_launch.setdefault(name, []).append(ms(a, b))
    kms = {k: sum(v) / len(v) for k, v in per_launch.items()}      # mean ms per launch

    # GLA decode kernel alone (the graph's GLA half also holds the gate launches and advance_step)
    # the product form: raw q / k / gk rows, feature map + gate inside the kernel (f2)
    gq, gk_, gv_ = randn(B, H, D), randn(B, Hkv, D), randn(B, Hkv, D)
    gg_ = randn(B, Hkv, D)
    og_ = torch.empty(B, H, D, device=dev, dtype=bf)
    gw = torch.ones(D, device=dev, dtype=bf)
    g_k = torch.cuda.CUDAGraph()
    for l in range(N_GLA_LAYERS):
        ops.gla_decode(gq, gk_, gv_, gg_, gla_state[l], idx_buf, out=og_, gnorm_weight=gw, gate_normalizer=16.0)
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g_k, stream=side):
            for l in range(N_GLA_LAYERS):
                ops.gla_decode(gq, gk_, gv_, gg_, gla_state[l], idx_buf, out=og_, gnorm_weight=gw,
                               gate_normalizer=16.0)
    torch.cuda.synchronize()
    a, b = ev(), ev()
    g_k.replay(); a.record()
    for _ in range(20):
        g_k.replay()
    b.record(); torch.cuda.synchronize()
    t_gdk_us = ms(a, b) * 1e3 / (20 * N_GLA_LAYERS)

    peaks = load_peaks()

    # ---- algorithmic bytes / FLOPs per launch (per GPU-local heads)
    b2 = 2
    mean_ctx = Tp + (Sd + 1) / 2
    nk = min(mean_ctx, WINDOW_LEFT + 1)
    swa_dec_bytes = B * (Hkv * 2 * D * b2 * nk + H * 2 * D * b2 + Hkv * 2 * D * b2 + 4 * (nk / bs) + 8)
    gla_dec_bytes = B * (Hkv * 2 * D * D * b2 + H * 2 * D * b2 + Hkv * D * (2 * b2 + 2))
    swa_pf_flops = B * sum(H * 4 * D * min(i + 1, WINDOW_LEFT + 1) for i in range(Tp))
    gla_pf_bytes = T * (H * 2 * D * b2 + Hkv * D * (2 * b2 + 2))
    gla_pf_flops = T * H * (4 * D * D + 2 * 64 * 2 * D)
    kvw_bytes = T * 2 * Hkv * D * b2 * 2
    us_dswa = t_dswa * 1e3 / (Sd * N_SWA_LAYERS)

    def K(share, bound, work, seconds, peak, unit, **kw):
        ach = work / seconds / (1e9 if unit == "GB/s" else 1e12)
        d = {"share_of_step": share, "bound": bound, "achieved": ach, "peak": peak, "unit": unit,
             "frac": ach / peak}
        d.update(kw)
        return d
    step_ms = step_s * 1e3
    kernels = {
        "swa_decode": K(t_dswa / step_ms, "hbm", swa_dec_bytes, us_dswa * 1e-6, peaks["hbm_gbs"], "GB/s",
                        us_per_layer_step=us_dswa,
                        what="one fused launch per layer-step: paged-KV write + split-K decode (+ combine)"),
        "gla_decode": K(t_dgla / step_ms, "hbm", gla_dec_bytes, t_gdk_us * 1e-6, peaks["hbm_gbs"], "GB/s",
                        us_per_layer_step_kernel=t_gdk_us,
                        us_per_layer_step_module=t_dgla * 1e3 / (Sd * N_GLA_LAYERS),
                        what="kernel = feature map + gate + recurrent update + g_norm epilogue on raw projection rows "
                             "(timed alone, graph of 14 launches); module = GatedLinearAttention.forward as "
                             "replayed in the step"),
        "swa_prefill": K(t_pswa / step_ms, "tensor", swa_pf_flops, kms["swa_prefill"] * 1e-3,
                         peaks["bf16_tflops_sustained"], "TFLOP/s", ms_per_layer_kernel=kms["swa_prefill"],
                         ms_per_layer_module=t_pswa / N_SWA_LAYERS,
                         frac_of_burst_bf16=swa_pf_flops / (kms["swa_prefill"] * 1e-3) / 1e12 / peaks["bf16_tflops"],
                         tok_s_per_layer=T / (t_pswa * 1e-3 / N_SWA_LAYERS)),
        "gla_prefill": K(t_pgla / step_ms, "hbm", gla_pf_bytes, kms["gla_prefill"] * 1e-3, peaks["hbm_gbs"], "GB/s",
                         ms_per_layer_kernel=kms["gla_prefill"], ms_per_layer_module=t_pgla / N_GLA_LAYERS,
                         tensor_tflops=gla_pf_flops / (kms["gla_prefill"] * 1e-3) / 1e12,
                         tok_s_per_layer=T / (t_pgla * 1e-3 / N_GLA_LAYERS),
                         what="ONE launch per layer on the raw projection rows: relu(q), relu(k), the gate "
                              "logsigmoid(gk)/16 looked up by the gate warps in a 64 K-entry table, g_norm epilogue"),
        "kv_write": K(kms["kv_write"] * N_SWA_LAYERS / step_ms, "hbm", kvw_bytes, kms["kv_write"] * 1e-3,
                      peaks["hbm_gbs"], "GB/s", ms_per_layer_kernel=kms["kv_write"]),
    }
    # DRAM bytes per launch from the committed `ncu --set full` captures (profiles/traffic.json)
    traffic = {}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f)
    except OSError:
        pass
    for name, kk in kernels.items():
        kk["traffic"] = traffic.get(name, {}).get("bytes_per_launch")
    dom = max(("swa_decode", "gla_decode", "swa_prefill", "gla_prefill"), key=lambda x: kernels[x]["share_of_step"])
    roofline = {"kernel": dom, "bound": kernels[dom]["bound"], "achieved": kernels[dom]["achieved"],
                "peak": kernels[dom]["peak"], "unit": kernels[dom]["unit"], "frac": kernels[dom]["frac"],
                "traffic": traffic.get(dom, {}).get("bytes_per_launch"),
                "traffic_source": traffic.get(dom, {}).get("source"),
                "algorithmic_bytes_per_launch": swa_dec_bytes if dom == "swa_decode" else None,
                "peak_source": f"{peaks['src']}, HBM copy bandwidth / sustained bf16",
                "note": "launch duration = CUDA events around the graph replays inside the timed region"}

    # ---- e2e: the same step through the same plugin calls from HOST buffers (pinned), with the
    # host-side work of a serving step inside the timed region: slot manager, metadata builder
    e2e = None
    if not args.no_e2e:
        dev_inputs = [swa_q, swa_k, swa_v, gla_qkv, gla_gk, dq, dk, dv, dg_qkv, dg_gk]
        host_inputs = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in dev_inputs]
        for hbuf, dbuf in zip(host_inputs, dev_inputs):
            hbuf.copy_(dbuf)
        host_out = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in (dout_swa, dout_gla)]
        meta_bytes = [0]

        def md_bytes(md):
            return sum(t.numel() * t.element_size() for t in (
                md.slot_mapping, md.seq_lens_tensor, md.query_start_loc, md.seq_start_loc,
                md.context_lens_tensor, md.block_tables) if t is not None)
        h2d = sum(t.numel() * t.element_size() for t in dev_inputs)
        d2h = sum(t.numel() * t.element_size() for t in host_out)
        # Copy pipeline (every byte of every timed step is copied inside the timed region):
        #  * decode inputs (small): main stream, start of the step;
        #  * GLA prefill inputs (half of the bytes): copy stream, underneath this step's SWA prefill,
        #    awaited on the device just before the GLA prefill;
        #  * SWA prefill inputs: double-buffered on the device -- the NEXT step's rows go out on the
        #    copy stream behind the GLA inputs, underneath this step's compute (what a serving loop
        #    does with the next batch); the first timed step has nothing to ride on and copies its
        #    own at its start, the last one prefetches nothing.
        swa_sets = [(swa_q, swa_k, swa_v), tuple(torch.empty_like(t) for t in (swa_q, swa_k, swa_v))]
        copy_s = torch.cuda.Stream()
        main_s = torch.cuda.current_stream()
        swa_ready = [None, None]                      # event per device buffer set

        def copy_swa(which, stream):
            with torch.cuda.stream(stream):
                for dbuf, hbuf in zip(swa_sets[which], host_inputs[:3]):
                    dbuf.copy_(hbuf, non_blocking=True)
                evt = torch.cuda.Event()
                evt.record(stream)
            swa_ready[which] = evt

        dbg = os.environ.get("HYMETA_BENCH_DEBUG") == "1"
        dbg_rec, dbg_copy = [], []

        def e2e_step(s, prefetch_next):
            th = [time.perf_counter()]
            cur = s & 1
            if swa_ready[cur] is None:                # not prefetched: copy now, on the critical path
                copy_swa(cur, main_s)
            for i in range(5, len(dev_inputs)):       # decode inputs
                dev_inputs[i].copy_(host_inputs[i], non_blocking=True)
            th.append(time.perf_counter())
            idx = admit_requests()                    # host: slot manager (release + admit 64 requests)
            th.append(time.perf_counter())
            mp, mdd = build_prefill_md(), build_decode_md()    # host: metadata builder + H2D of its tensors
            th.append(time.perf_counter())
            meta_bytes[0] = md_bytes(mp) + md_bytes(mdd) + idx.numel() * 4
            # the bulk copies are enqueued AFTER the step's small ones (decode inputs, metadata, slot
            # indices): all host->device copies share one DMA queue, and a 4 MB metadata copy enqueued
            # behind 10 GB of activations held the step's first kernel back by ~180 ms
            copy_s.wait_stream(main_s)                # the previous step is done with the shared buffers
            with torch.cuda.stream(copy_s):
                c0 = ev(); c0.record(copy_s)
                for i in (3, 4):                      # gla_qkv, gla_gk
                    dev_inputs[i].copy_(host_inputs[i], non_blocking=True)
                gla_ready = ev()
                gla_ready.record(copy_s)
            if prefetch_next:
                copy_swa(cur ^ 1, copy_s)
            if dbg:
                c2 = ev(); c2.record(copy_s)
                dbg_copy.append((c0, gla_ready, c2))
            main_s.wait_event(swa_ready[cur])
            swa_ready[cur] = None
            one_step(rec=dbg_rec if dbg else None, md=mp, idx=idx,
                     before_gla=lambda: main_s.wait_event(gla_ready), fresh_md_d=mdd, swa_in=swa_sets[cur])
            th.append(time.perf_counter())
            host_out[0].copy_(dout_swa, non_blocking=True)
            host_out[1].copy_(dout_gla, non_blocking=True)
            th.append(time.perf_counter())
            if dbg and rank == 0:
                sys.stderr.write("e2e host ms: copies %.1f  slot mgr %.1f  metadata %.1f  launches %.1f  d2h %.1f\n" %
                                 tuple((b - a) * 1e3 for a, b in zip(th, th[1:])))
        e2e_step(0, False)                            # warm-up: prefetches nothing for the timed steps
        barrier()
        k = max(1, args.e2e_steps)
        s2, p2 = ev(), ev()
        tw0 = time.perf_counter()
        s2.record()
        for s in range(k):
            e2e_step(s + 1, s + 1 < k)
        p2.record()
        barrier()
        tw1 = time.perf_counter()
        el = s2.elapsed_time(p2) * 1e-3
        if dbg and rank == 0:
            gb = lambda idxs: sum(dev_inputs[i].numel() * 2 for i in idxs) / 1e9
            for (e, marks), (c0, c1, c2) in zip(dbg_rec, dbg_copy):
                sys.stderr.write("e2e copy stream: starts %.1f ms after the step's first kernel; GLA inputs %.1f ms "
                                 "(%.1f GB/s); next SWA inputs %.1f ms (%.1f GB/s)\n" %
                                 (ms(e[0], c0), ms(c0, c1), gb((3, 4)) / max(ms(c0, c1), 1e-3) * 1e3, ms(c1, c2),
                                  gb((0, 1, 2)) / max(ms(c1, c2), 1e-3) * 1e3))
            for e, marks in dbg_rec:
                sys.stderr.write("e2e gpu ms: swa prefill %.1f  gla prefill (incl. input wait) %.1f  decode %.1f "
                                 "(swa %.1f gla %.1f)\n" % (ms(e[0], e[1]), ms(e[1], e[2]), ms(e[2], e[3]),
                                                           sum(ms(a, b) for a, b, c in marks),
                                                           sum(ms(b, c) for a, b, c in marks)))
        if world > 1:
            t = torch.tensor([el], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            el = t.item()
        e2e = {"value": total_tokens / (el / k), "unit": "tok/s", "h2d_bytes_per_step": h2d + meta_bytes[0],
               "d2h_bytes_per_step": d2h, "steps": k, "ms_per_step": el / k * 1e3,
               "wall_ms_per_step": (tw1 - tw0) / k * 1e3,
               "note": "per step: activations (one layer-set of q/k/v and qkv/gk rows, decode inputs) copied "
                       "from pinned host memory inside the timed region (GLA inputs on a copy stream under the SWA "
                       "prefill; the next step's SWA inputs double-buffered under this step's compute, the "
                       "first timed step's on its critical path); slot "
                       "manager admits/releases the requests; MetaAttentionMetadataBuilder builds prefill and "
                       "decode metadata on the host and ships it; MetaAttentionImpl.forward / "
                       "GatedLinearAttention.forward per layer; last decode outputs read back"}

    cpu = None
    if rank == 0 and n == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(args, os.cpu_count() or 1)

    # ---- extra legs, outside every timed region above
    extra = {}
    if not args.no_extra:
        del swa_q, swa_k, swa_v, gla_qkv, gla_gk, out_swa, kv_caches
        if not args.no_e2e:
            del dev_inputs, host_inputs, swa_sets
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_legs
        tp_group = None
        if world > 1:
            for d in range(plan.dp):                   # one group per head-shard set
                grp = dist.new_group([d * plan.tp + r for r in range(plan.tp)])
                if d == plan.dp_rank:
                    tp_group = grp
        try:
            extra["configs"] = bench_legs.extra_configs(dev, world, rank, tp_group, plan)
        except Exception as e:
            extra["configs"] = {"error": f"{type(e).__name__}: {str(e)[:300]}"}
        if n == 1:
            try:
                extra["vs_live"] = bench_legs.vs_live(dev)
            except Exception as e:
                extra["vs_live"] = {"error": f"{type(e).__name__}: {str(e)[:300]}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "tok/s", "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args, n), "e2e": e2e, "gpu_launches": int(gpu_launches),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "detail": {"prefill_tok_s": args.batch * n * Tp / ((t_pswa + t_pgla) * 1e-3),
                       "decode_us_per_step": t_dec * 1e3 / Sd, "kernels": kernels,
                       "launches_per_decode_step": launches_swa_step + launches_gla_step,
                       "tokens_per_gpu_equiv": tokens_per_gpu,
                       "per_gpu": {"q_heads": H, "kv_heads": Hkv, "requests": B}, **extra},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
