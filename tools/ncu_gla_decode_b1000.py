"""ncu target: a few eager GLA decode launches at B=1000 (config 4 shape), compact state layout."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_hymeta_b200 import ops
dev, B, H, Hk, D = "cuda", 1000, 28, 4, 128
q = torch.randn(B, H, D, device=dev).bfloat16().relu()
k = torch.randn(B, Hk, D, device=dev).bfloat16().relu()
v = torch.randn(B, Hk, D, device=dev).bfloat16()
g = (-torch.rand(B, Hk, D, device=dev) * 0.1).bfloat16()
state = torch.zeros(B, Hk, D, D, device=dev, dtype=torch.bfloat16)
slots = torch.randperm(B, device=dev).to(torch.int32)
out = torch.empty(B, H, D, device=dev, dtype=torch.bfloat16)
for _ in range(4):
    ops.gla_decode(q, k, v, g, state, slots, out=out)
torch.cuda.synchronize()
