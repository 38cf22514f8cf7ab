#!/usr/bin/env python3
"""Debug aid: run the T=4096 single-chain GLA prefill of tests/test_gpu_gla_prefill.py through a
given build of the library (e.g. a GP_EXP variant from ab_swa_prefill.py --src gla_prefill) several
times and print the error against the fp32 oracle."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vllm_hymeta_b200 import _lib
if len(sys.argv) > 1:
    _lib.LIB_PATH = sys.argv[1]
from oracle import gla as ogla
from vllm_hymeta_b200 import ops

H, Hkv, D, T = 28, 4, 128, 4096
q, k, v, g = ogla.synth_gla_inputs(T, H, Hkv, D, "70b", seed=11)
ref_o, ref_S = ogla.gla_chunked(q, k, v, g, chunk=16)
dq, dk, dv, dg = (x.cuda() for x in (q, k, v, g))
i32 = lambda x: torch.tensor(x, dtype=torch.int32, device="cuda")
nbad = 0
for rep in range(40):
    seg_len = 512
    st = torch.zeros(2, Hkv, D, D, dtype=torch.bfloat16, device="cuda")
    o = ops.gla_prefill_varlen(dq, dk, dv, dg, st, i32([0, T]), i32([0]), i32([0]), seg_len=seg_len)
    d = (o.float().cpu() - ref_o.float()).abs()
    tok = d.amax(dim=(1, 2))
    bad = (tok > 2 ** -7 * ref_o.abs().max()).nonzero().flatten()
    nbad += int(bad.numel() > 0)
    if bad.numel() and nbad <= 3:
        tol = 2 ** -7 * ref_o.abs().max()
        for tk in bad[:12].tolist():
            hd = (d[tk].amax(dim=1) > tol).nonzero().flatten().tolist()
            cols = (d[tk] > tol).sum(dim=1)
            print(f"   token {tk} (step {tk // 16} t {tk % 16}): heads {hd} bad cols per head {[int(cols[h]) for h in hd]} "
                  f"got0 {o[tk, hd[0], :4].float().tolist()} ref0 {ref_o[tk, hd[0], :4].tolist()}")
    if bad.numel() == 0 and rep < 39:
        continue
    print(f"[{nbad} bad runs so far] {os.path.basename(_lib.LIB_PATH)} rep {rep} seg {seg_len}: max|d| {d.max().item():.4g} (tol {2**-7*ref_o.abs().max().item():.4g}) "
          f"bad tokens {bad.numel()} first {bad[:6].tolist()} last {bad[-3:].tolist()} heads {d.amax(dim=(0,2)).argmax().item()}", flush=True)
