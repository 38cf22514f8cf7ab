"""Which arithmetic does the INSTALLED vLLM's rotary_embedding op use?  (The reference pins vLLM
0.10.0, whose kernel rounds to bf16 after every product and sum; the image has 0.22.)  Prints the
fraction of output elements each candidate reproduces bit for bit."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_hymeta_b200 import ops  # noqa: E402
import vllm._custom_ops as vops  # noqa: E402

D, T, H = 128, 4096, 8
dev = "cuda"
gen = torch.Generator().manual_seed(0)
q = torch.randn(T, H * D, generator=gen).bfloat16().to(dev)
pos = torch.randint(0, 131072, (T,), generator=gen).to(dev)
table = ops.rope_cos_sin_cache(D, 131072, 1e6).to(dev)
ref = q.clone()
vops.rotary_embedding(pos, ref, None, D, table, True)
ref = ref.view(T, H, D)
x, y = q.view(T, H, D)[..., :64], q.view(T, H, D)[..., 64:]
cs = table[pos]
cos, sin = cs[:, None, :64], cs[:, None, 64:]
cands = {}
cands["bf16 after every op (vLLM 0.10.0 c10::BFloat16 operators)"] = torch.cat([x * cos - y * sin, y * cos + x * sin], -1)
xf, yf, cf, sf = x.float(), y.float(), cos.float(), sin.float()
cands["fp32 mul/sub, one rounding"] = torch.cat([xf * cf - yf * sf, yf * cf + xf * sf], -1).bfloat16()
cands["fp32 fma(x,cos,-(y*sin)), one rounding"] = torch.cat(
    [torch.addcmul(-(yf * sf), xf, cf), torch.addcmul(xf * sf, yf, cf)], -1).bfloat16()
cands["fp32 fma(-y,sin,x*cos), one rounding"] = torch.cat(
    [torch.addcmul(xf * cf, -yf, sf), torch.addcmul(yf * cf, xf, sf)], -1).bfloat16()
ours = q.clone()
ops.rope_kv_write_paged(ours.view(T, H, D), ours.view(T, H, D)[:, :1].clone(), None, table, pos)
cands["hymeta_rope_kv_write_paged"] = ours.view(T, H, D)
for name, c in cands.items():
    same = (c == ref).float().mean().item()
    md = (c.float() - ref.float()).abs().max().item()
    print(f"{same:9.6f} identical, max|d| {md:.4g}  <- {name}")
