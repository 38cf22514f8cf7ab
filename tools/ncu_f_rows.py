#!/usr/bin/env python3
"""ncu target for the f-row kernels: one large launch each of the fused SWA prologue (f3), the
branch merge (f1) and the g_norm decode epilogue (f2).  tools/profile.sh-style use:
  ncu --set full --clock-control none --import-source on -k regex:<kernel> -s 1 -c 1 -o gpurun_out/prof_<kernel> python tools/ncu_f_rows.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from vllm_hymeta_b200 import ops

dev, D, H, Hkv, bs = "cuda", 128, 28, 4, 256
T = 524288
nb = T // bs
kv = torch.zeros(2, nb, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
qkv = torch.randn(T, (H + 2 * Hkv) * D, device=dev).bfloat16()
q, k, v = (x.view(T, -1, D) for x in qkv.split([H * D, Hkv * D, Hkv * D], dim=-1))
table = ops.rope_cos_sin_cache(D, 16384, 1e6, device=dev)
pos = torch.randint(0, 16000, (T,), device=dev)
slots = torch.randperm(nb * bs, device=dev)[:T].to(torch.int64)
for _ in range(3):
    ops.rope_kv_write_paged(q, k, v, table, pos, 128, kv[0], kv[1], slots)
Tm, hidden = 65536, 3584
xa, xl = (torch.randn(Tm, hidden, device=dev).bfloat16() for _ in range(2))
w1, w2 = torch.ones(hidden, device=dev), torch.ones(hidden, device=dev)
out = torch.empty_like(xa)
for _ in range(3):
    ops.branch_norm_merge(xa, xl, w1, w2, 1e-6, out=out)
B = 1000
state = torch.zeros(B, Hkv, D, D, device=dev, dtype=torch.bfloat16)
dq = torch.randn(B, H, D, device=dev).bfloat16()
dk, dv = (torch.randn(B, Hkv, D, device=dev).bfloat16() for _ in range(2))
dg = (-torch.rand(B, Hkv, D, device=dev) * 0.1).bfloat16()
sl = torch.arange(B, device=dev, dtype=torch.int32)
gw = torch.ones(D, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    ops.gla_decode(dq, dk, dv, dg, state, sl, gnorm_weight=gw)
torch.cuda.synchronize()
print("ok")
