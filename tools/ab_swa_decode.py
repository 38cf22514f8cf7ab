"""A/B of the SWA decode launch structure at config 2 (B=64, ctx 8448, 14 layers in one CUDA graph):
(a) kv_write + split + combine, (b) fused write + split + combine kernel, (c) fully fused."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_hymeta_b200 import ops
dev = "cuda"
B = int(os.environ.get("AB_B", "64"))
H, Hkv, D, bs, ctx = 28, 4, 128, 256, 8448
nblk = (ctx + bs) // bs
L = max(2, min(14, int(10e9 / (B * nblk * bs * Hkv * D * 4))))
nb = B * nblk
kv = [torch.randn(2, nb, bs, Hkv, D, device=dev, dtype=torch.bfloat16) for _ in range(L)]
tables = torch.randperm(nb, device=dev).to(torch.int32).view(B, nblk)
lens = torch.full((B,), ctx, dtype=torch.int32, device=dev)
q = torch.randn(B, H, D, device=dev, dtype=torch.bfloat16)
nk = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
nv = torch.randn(B, Hkv, D, device=dev, dtype=torch.bfloat16)
pos = lens.long() - 1
slot = (tables.long().gather(1, (pos // bs).unsqueeze(1)).squeeze(1) * bs + pos % bs)
out = torch.empty(B, H, D, device=dev, dtype=torch.bfloat16)
def va():
    for l in range(L):
        ops.kv_write_paged(nk, nv, kv[l][0], kv[l][1], slot)
        ops.swa_meta_decode_paged(q, kv[l][0], kv[l][1], tables, lens, None, None, 4096, out=out, fused_combine=False)
def vb():
    for l in range(L):
        ops.swa_meta_decode_paged(q, kv[l][0], kv[l][1], tables, lens, None, None, 4096, out=out, new_k=nk, new_v=nv, fused_combine=False)
def vc():
    for l in range(L):
        ops.swa_meta_decode_paged(q, kv[l][0], kv[l][1], tables, lens, None, None, 4096, out=out, new_k=nk, new_v=nv)
variants = (("write+split+combine", va), ("fusedwrite-split+combine", vb), ("fully fused", vc))
if os.environ.get("AB_ONLY"):
    variants = variants[1:2]
for name, fn in variants:
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph(); st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        with torch.cuda.graph(g, stream=st):
            fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(7):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); g.replay(); e.record(); torch.cuda.synchronize(); ts.append(s.elapsed_time(e))
    ts.sort()
    print(f"{name:28s} {ts[len(ts)//2]*1e3/L:8.1f} us per layer-step  B={B} L={L} TPS={os.environ.get('HYMETA_SWA_DEC_TPS','model')}")
