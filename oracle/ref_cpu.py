"""CPU timing legs of bench.py.  TEST / BENCH INFRASTRUCTURE -- see oracle/__init__.py.

Two things live here:

* ``port_stack_sample``: the oracle port of the hot-path ARITHMETIC (no projections) at the
  config-2 shape -- one 7B GLA layer + one SWA layer, B = 1, prefill ``T`` tokens (``T`` >= the
  4096-key window, so the window actually binds) + ``decode`` single-token steps.
* ``reference_config1``: BASELINE.json configs[0] -- the reference's OWN ``hf_7B_model``
  modules (``GatedLinearAttention`` + ``FlashAttention``, unmodified, loaded from ``baseline/_ref``)
  on CPU at the 7B shape, B = 1, T = 2048, with only the GPU-only third-party symbols replaced in
  the modules' namespaces: fla's Triton ops -> fla's own ``naive_recurrent_gla``,
  fla's Triton RMSNorm -> the reference's torch ``hf_7B_model/norms.py``, flash-attn's varlen call
  -> a dense masked softmax following ``naive_metatoken_flash_attn``.  These modules include the
  q/k/v/o projections, which the kernels of this repository do not replace; the figure is reported
  as its own entry and never mixed into the hot-path ratio.
"""
from __future__ import annotations

import os
import sys
import time
from typing import Dict, Optional

import torch

from . import gla as ogla
from . import swa as oswa

H, HKV, D, W_LEFT = 28, 4, 128, 4096


def swa_dense_blocked(q, k, v, window_left: int, block: int = 1024) -> torch.Tensor:
    """``swa_meta_dense`` evaluated per block of queries over the keys that block can see (the dense
    [T, T] score matrix of a long prompt does not fit in host memory)."""
    T = q.shape[0]
    outs = []
    for s in range(0, T, block):
        e = min(T, s + block)
        lo = 0 if window_left < 0 else max(0, s - window_left)
        o, _ = oswa.swa_meta_dense(q[s:e], k[lo:e], v[lo:e], None, None, None, window_left)
        outs.append(o)
    return torch.cat(outs, 0)


def gla_decode_batched(q, k, v, g, S):
    """One recurrent step for a batch, fp32, reference layout (one state per q-head after repeat_kv;
    my_fused_recurrent.py:58-80 / fla naive.py:33-40).  q [B,H,D]; k,v,g [B,Hkv,D]; S [B,H,D,D] in place."""
    Hq = q.shape[1]
    rep = lambda x: x.repeat_interleave(Hq // x.shape[1], dim=1)
    k, v, g = rep(k), rep(v), rep(g)
    S.mul_(g.exp().unsqueeze(-1)).add_(k.unsqueeze(-1) * v.unsqueeze(-2))
    return torch.einsum("bhd,bhde->bhe", q * D ** -0.5, S)


def swa_decode_batched(q, K, V):
    """One decode token per request over its visible keys, dense softmax, GQA without repeat.
    q [B,H,D]; K,V [B,n,Hkv,D] (the window already cut)."""
    B, Hq, _ = q.shape
    Hkv = K.shape[2]
    qg = q.view(B, Hkv, Hq // Hkv, D) * D ** -0.5
    p = torch.softmax(torch.einsum("bhgd,bnhd->bhgn", qg, K), dim=-1)
    return torch.einsum("bhgn,bnhd->bhgd", p, V).reshape(B, Hq, D)


def port_stack_sample(T: int, decode: int, threads: int, seed: int = 0, decode_batch: int = 16) -> Dict[str, float]:
    """Seconds for one GLA layer and one SWA layer of the 7B stack on the host cores (fp32): prefill
    of ONE ``T``-token prompt, then ``decode`` steps of a batch of ``decode_batch`` requests at
    context ``T`` + step (batched tensor ops, as the reference's torch path batches them)."""
    torch.set_num_threads(threads)
    q, k, v, g = ogla.synth_gla_inputs(T, H, HKV, D, "7b", seed=seed, dtype=torch.float32)
    sq, sk, sv, _, _ = oswa.synth_swa_inputs(T, H, HKV, D, seed=seed + 1, dtype=torch.float32)
    Bd = decode_batch
    gen = torch.Generator().manual_seed(seed + 2)
    dq, dk, dv = (torch.randn(decode, Bd, h, D, generator=gen) for h in (H, HKV, HKV))
    dg = torch.nn.functional.logsigmoid(torch.randn(decode, Bd, HKV, D, generator=gen)) / 16
    Kc = torch.randn(Bd, T + decode, HKV, D, generator=gen)
    Vc = torch.randn(Bd, T + decode, HKV, D, generator=gen)
    t0 = time.perf_counter()
    o, S = ogla.gla_recurrent(q, k, v, g)
    t1 = time.perf_counter()
    Sb = S.unsqueeze(0).repeat(Bd, 1, 1, 1).contiguous()
    t1b = time.perf_counter()
    for t in range(decode):
        gla_decode_batched(dq[t].relu(), dk[t].relu(), dv[t], dg[t], Sb)
    t2 = time.perf_counter()
    swa_dense_blocked(sq, sk, sv, W_LEFT)
    t3 = time.perf_counter()
    for t in range(decode):
        L = T + t + 1
        lo = max(0, L - 1 - W_LEFT)
        Kc[:, L - 1], Vc[:, L - 1] = dk[t], dv[t]                  # the step's KV write
        swa_decode_batched(dq[t], Kc[:, lo:L], Vc[:, lo:L])
    t4 = time.perf_counter()
    return {"gla_prefill_s": t1 - t0, "gla_decode_s": t2 - t1b, "swa_prefill_s": t3 - t2,
            "swa_decode_s": t4 - t3, "total_s": t4 - t0, "prefill_tokens": T,
            "decode_tokens": decode * Bd}


def stack_tok_s(sample: Dict[str, float], prefill_len: int, decode_steps: int, layer_pairs: int = 14) -> float:
    """Tokens/s of the 14 + 14-layer stack at the workload's prefill : decode mix, from the
    per-token prefill and decode costs of a one-layer-pair sample (the CPU cost is linear in the
    batch, so B = 1 stands for any batch)."""
    per_pf = (sample["gla_prefill_s"] + sample["swa_prefill_s"]) / sample["prefill_tokens"]
    per_dc = (sample["gla_decode_s"] + sample["swa_decode_s"]) / max(1, sample["decode_tokens"])
    return (prefill_len + decode_steps) / ((prefill_len * per_pf + decode_steps * per_dc) * layer_pairs)


# --------------------------------------------------------------------------------------------------
def _ref_root() -> Optional[str]:
    from . import install_ref
    return install_ref.TARGET if install_ref.installed() else None


def reference_config1(threads: int, T: int = 2048, reps: int = 2) -> Optional[Dict[str, object]]:
    """BASELINE.json configs[0] through the reference's own modules; None if ``baseline/_ref`` (or
    fla's naive op) is unavailable."""
    root = _ref_root()
    if root is None:
        return None
    try:
        from fla.ops.gla.naive import naive_recurrent_gla
    except Exception:
        return None
    if root not in sys.path:
        sys.path.insert(0, root)
    try:
        import hf_7B_model.gla_attention as mg
        import hf_7B_model.norms as norms
        import hf_7B_model.window_attention as mw
    except Exception:
        return None
    torch.set_num_threads(threads)
    tr = lambda x: x.transpose(1, 2)

    def gla_shim(q, k, v, gk, initial_state=None, output_final_state=False, **kw):
        o, h = naive_recurrent_gla(tr(q), tr(k), tr(v), tr(gk), initial_state=initial_state,
                                   output_final_state=output_final_state)
        return tr(o), h

    mg.chunk_gla = mg.fused_chunk_gla = mg.fused_recurrent_gla = gla_shim
    mg.RMSNorm = norms.RMSNorm

    def upad(q, k, v, mask, q_len, unpad_fn=None):
        B, L = k.shape[:2]
        cu = torch.arange(0, (B + 1) * L, L, dtype=torch.int32)
        return (q.reshape(-1, *q.shape[2:]), k.reshape(-1, *k.shape[2:]), v.reshape(-1, *v.shape[2:]),
                None, (cu, cu), (q_len, L))

    def varlen(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p=0.0,
               softmax_scale=None, causal=True, window_size=(-1, -1), **kw):
        outs = [swa_dense_blocked(q[int(a):int(b)], k[int(a):int(b)], v[int(a):int(b)], int(window_size[0]))
                for a, b in zip(cu_seqlens_q[:-1], cu_seqlens_q[1:])]
        return torch.cat(outs, 0).to(q.dtype)

    mw._upad_input = upad
    mw.unpad_input = None
    mw.flash_attn_varlen_func = varlen
    mw.pad_input = lambda x, idx, B, L: x.view(B, L, *x.shape[1:])
    hidden = H * D
    with torch.no_grad():
        torch.manual_seed(0)
        gla = mg.GatedLinearAttention(mode="chunk", hidden_size=hidden, num_heads=H,
                                      num_key_value_heads=HKV, layer_idx=0).eval()
        swa = mw.FlashAttention(hidden_size=hidden, num_heads=H, num_key_value_heads=HKV,
                                max_position_embeddings=131072, sliding_window=W_LEFT, rope_theta=1e6,
                                layer_idx=1).eval()
        x = torch.randn(1, T, hidden)
        pos = torch.arange(T)[None]
        best_g = best_s = float("inf")
        for _ in range(reps):
            t0 = time.perf_counter()
            gla(x)
            t1 = time.perf_counter()
            swa(x, position_ids=pos)
            t2 = time.perf_counter()
            best_g, best_s = min(best_g, t1 - t0), min(best_s, t2 - t1)
    return {"config": "BASELINE.json configs[0]: hf_7B_model GatedLinearAttention + FlashAttention "
                      f"(unmodified modules from baseline/_ref, GPU-only third-party symbols shimmed), "
                      f"B=1, T={T}, fp32, random init, projections included",
            "gla_layer_s": best_g, "swa_layer_s": best_s, "tok_s_per_layer_pair": T / (best_g + best_s),
            "cores": threads, "kind": "reference"}


if __name__ == "__main__":
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    n = os.cpu_count() or 1
    print(reference_config1(n))
    smp = port_stack_sample(4608, 16, n)
    print(smp, stack_tok_s(smp, 8192, 512))
