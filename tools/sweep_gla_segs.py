#!/usr/bin/env python3
"""Sweep the segment count of the sequence-parallel GLA prefill for few long chains (config 5)."""
import os, sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from vllm_hymeta_b200 import ops

D = 128
def run(n_req, hkv, T, segs):
    dev, h = "cuda", 7 * hkv
    q = torch.randn(n_req * T, h, D, device=dev).bfloat16().relu()
    k = torch.randn(n_req * T, hkv, D, device=dev).bfloat16().relu()
    v = torch.randn(n_req * T, hkv, D, device=dev).bfloat16()
    g = (torch.nn.functional.logsigmoid(torch.randn(n_req * T, hkv, D, device=dev)) / 16).bfloat16()
    state = torch.zeros(n_req, hkv, D, D, device=dev, dtype=torch.bfloat16)
    cu = (torch.arange(n_req + 1, device=dev) * T).to(torch.int32)
    sl = torch.arange(n_req, device=dev, dtype=torch.int32)
    md = torch.zeros(n_req, device=dev, dtype=torch.int32)
    out = torch.empty(n_req * T, h, D, device=dev, dtype=torch.bfloat16)
    res = []
    for ns in segs:
        seg = 0 if ns <= 1 else -(-(-(-T // ns)) // 16) * 16
        f = lambda: ops.gla_prefill_varlen(q, k, v, g, state, cu, sl, md, max_seqlen=T, out=out, seg_len=seg)
        for _ in range(2): f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): f()
        e1.record(); torch.cuda.synchronize()
        res.append((ns, e0.elapsed_time(e1) / 5))
    auto = ops.gla_prefill_seg_len(n_req * hkv, T)
    best = min(res, key=lambda x: x[1])
    print(f"chains={n_req*hkv} T={T} auto seg_len={auto} (n_segs={-(-T//auto) if auto else 1}): " +
          " ".join(f"{ns}:{ms:.3f}" for ns, ms in res) + f"  best n_segs={best[0]} {best[1]:.3f} ms")

segs = [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 16, 18, 20, 24, 28, 32, 37, 48, 64, 74]
run(8, 4, 32768, segs)
run(8, 4, 8192, [s for s in segs if s <= 32])
run(8, 4, 131072, segs)
run(8, 1, 32768, segs + [96, 128])
run(4, 1, 32768, segs + [96, 128])
run(2, 4, 32768, segs)
run(16, 4, 8192, [s for s in segs if s <= 32])
