#!/usr/bin/env python3
"""GLA prefill at the config-2 shape (64 x 8192, 28/4 heads): one chain per CTA (seg_len 0) against the
balanced work lists (seg_len -1), plain and with the 7B relu + g_norm epilogue; checks that the two
forms produce the same bits.  HYMETA_B200_LIB=<A/B library> selects another build of the kernels."""
import sys, torch
sys.path.insert(0, "/root/repo")
from vllm_hymeta_b200 import ops
dev = "cuda"
N, T = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (64, 8192)   # requests x prompt length
H, Hkv, D = 28, 4, 128
q = torch.randn(N * T, H, D, device=dev).bfloat16()
k = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
v = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
g = (torch.nn.functional.logsigmoid(torch.randn(N * T, Hkv, D, device=dev)) / 16).bfloat16()
if len(sys.argv) > 3 and sys.argv[3] == "strided":   # q / k / v as views of one qkv projection output, as the module passes them
    qkv = torch.cat([q.view(N * T, -1), k.view(N * T, -1), v.view(N * T, -1)], 1)
    q = qkv[:, :H * D].view(N * T, H, D)
    k = qkv[:, H * D:(H + Hkv) * D].view(N * T, Hkv, D)
    v = qkv[:, (H + Hkv) * D:].view(N * T, Hkv, D)
state = torch.zeros(N, Hkv, D, D, device=dev, dtype=torch.bfloat16)
cu = torch.arange(0, (N + 1) * T, T, device=dev, dtype=torch.int32)
slots = torch.arange(N, device=dev, dtype=torch.int32)
modes = torch.zeros(N, device=dev, dtype=torch.int32)
out = torch.empty_like(q)
gw = torch.ones(D, device=dev, dtype=torch.bfloat16)
outs = {}
graw = torch.randn(N * T, Hkv, D, device=dev).bfloat16()      # raw gk rows for the gate-table form
gfull = g
for name, kw in (("plain", {}), ("gn", dict(gnorm_weight=gw, relu_qk=True)),
                 ("fused", dict(gnorm_weight=gw, relu_qk=True, gate_table=ops.gla_gate_7b_table(16.0, q.device)))):
    g = graw if name == "fused" else gfull
    for seg in (0, -1):
        fn = lambda: ops.gla_prefill_varlen(q, k, v, g, state, cu, slots, modes, max_seqlen=T, seg_len=seg, out=out, **kw)
        for _ in range(2): fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(7):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        outs[(name, seg)] = out.clone()
        print(f"{name} seg_len={seg}: {ts[len(ts)//2]:.3f} ms (min {ts[0]:.3f})", flush=True)
    print(name, "equal:", torch.equal(outs[(name, 0)], outs[(name, -1)]))
