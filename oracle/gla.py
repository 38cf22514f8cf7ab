"""Oracle: gated linear attention (GLA).  TEST INFRASTRUCTURE -- see oracle/__init__.py.

Layouts follow the vLLM side of the reference: activations are token-major
``[T, heads, D]``; the recurrent state of one (request, head) is ``[d_k, d_v]``
row-major (vllm_hymeta/models/my_fused_recurrent.py:41-44).

Recurrence (my_fused_recurrent.py:58-80; fla/ops/gla/naive.py:33-40):

    S_t = exp(g_t)[:, None] * S_{t-1} + k_t (outer) v_t
    o_t = (q_t * D**-0.5)^T S_t

``k, v, g`` may carry fewer heads than ``q`` (GQA, un-repeated): q-head ``h``
then uses kv-head ``h // (H // Hkv)``, which is exactly what the reference's
``repeat_kv`` materialises (modeling_hymeta.py:607-609, gla_attention.py:264-267).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

INIT_ZERO = 0    # initial_state=None              (modeling_hymeta.py:489-490)
INIT_SLOT = 1    # initial_state=kv_cache[slot]    (modeling_hymeta.py:488)
INIT_META = 2    # initial_state=meta_linear_cache (modeling_hymeta.py:501-504)


def _expand_heads(x: torch.Tensor, H: int) -> torch.Tensor:
    """[T, Hkv, D] -> [T, H, D] the way repeat_kv does (interleaved repeat)."""
    Hkv = x.shape[1]
    if Hkv == H:
        return x
    assert H % Hkv == 0
    return x.repeat_interleave(H // Hkv, dim=1)


# --------------------------------------------------------------------------- a4
def gate_70b(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
             lower_bound: Optional[torch.Tensor] = None, layer_idx: int = 0,
             clamp_max: float = 0.95, num_heads: Optional[int] = None,
             head_start: int = 0, rule: str = "vllm"):
    """70B feature map + gate.  modeling_hymeta.py:611-629 (HF twin
    hf_model/linear_attention.py:143-157).

    q: [T,H,D] bf16, k: [T,Hk,D] bf16 (Hk = H after repeat_kv or Hkv before --
    the math is per element).  ``lower_bound``: full ``[hidden]`` row, sliced to
    the local heads ``head_start : head_start+H`` as at :622-625.
    ``rule``: "vllm" -> first-layer test is ``layer_idx % 7 == 0`` (:616),
    "hf" -> ``layer_idx == 0`` (hf_model/linear_attention.py:148).
    Returns (relu(q), k' in v.dtype, g fp32).
    """
    q = torch.relu(q)
    kf = k.float()
    first = (layer_idx % 7 == 0) if rule == "vllm" else (layer_idx == 0)
    if lower_bound is None or first:
        kf = kf.sigmoid()
        kf = torch.clamp(kf, max=clamp_max)
        g = (1 - kf).log()
    else:
        kf = kf.sigmoid()
        H = num_heads if num_heads is not None else lower_bound.numel() // k.shape[-1]
        lb = lower_bound.reshape(1, H, k.shape[-1])
        lb = lb[:, head_start:head_start + k.shape[1], :]
        g = 1 - (1 - lb) * kf
        kf, g = 1 - g, g.log()
    return q, kf.to(v.dtype), g


def gate_7b(q: torch.Tensor, k: torch.Tensor, gk: torch.Tensor, normalizer: float = 16.0):
    """7B feature map + gate.  model_for_7B/gla_attention.py:251-262
    (HF twin hf_7B_model/gla_attention.py:134-149).  Stays in the input dtype
    (bf16 in the reference), as the reference does."""
    return torch.relu(q), torch.relu(k), F.logsigmoid(gk) / normalizer


# --------------------------------------------------------------------------- a1
def gla_recurrent(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, g: torch.Tensor,
                  initial_state: Optional[torch.Tensor] = None,
                  scale: Optional[float] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """fp32 step recurrence for ONE sequence.

    q: [T,H,D]; k,v,g: [T,Hk,D] with H % Hk == 0; initial_state: [Hs,D,D]
    (Hs == H or Hs == Hk; a state shared by a GQA group is mathematically
    identical for all q-heads of the group -- SURVEY F7).
    Returns (o [T,H,D] fp32, final state [H,D,D] fp32).
    """
    T, H, D = q.shape
    if scale is None:
        scale = D ** -0.5
    qf = q.float() * scale
    kf = _expand_heads(k.float(), H)
    vf = _expand_heads(v.float(), H)
    gf = _expand_heads(g.float(), H)
    S = torch.zeros(H, D, v.shape[-1], dtype=torch.float32)
    if initial_state is not None:
        S = S + _expand_heads(initial_state.float().unsqueeze(0), H)[0]
    o = torch.empty(T, H, v.shape[-1], dtype=torch.float32)
    for t in range(T):
        S = S * gf[t].exp().unsqueeze(-1) + kf[t].unsqueeze(-1) * vf[t].unsqueeze(-2)
        o[t] = (qf[t].unsqueeze(-1) * S).sum(-2)
    return o, S


def gla_chunked(q, k, v, g, initial_state=None, scale=None, chunk: int = 64):
    """Same result as gla_recurrent in closed (chunk-parallel) form, fp64 inside;
    used as an independent cross-check of the recurrence and for long sequences."""
    T, H, D = q.shape
    if scale is None:
        scale = D ** -0.5
    qf = (q.double() * scale)
    kf = _expand_heads(k.double(), H)
    vf = _expand_heads(v.double(), H)
    gf = _expand_heads(g.double(), H)
    S = torch.zeros(H, D, v.shape[-1], dtype=torch.float64)
    if initial_state is not None:
        S = S + _expand_heads(initial_state.double().unsqueeze(0), H)[0]
    o = torch.empty(T, H, v.shape[-1], dtype=torch.float64)
    for s in range(0, T, chunk):
        e = min(T, s + chunk)
        b = gf[s:e].cumsum(0)                                   # [C,H,D]
        qc, kc, vc = qf[s:e], kf[s:e], vf[s:e]
        # inter-chunk
        o_c = torch.einsum("chd,hde->che", qc * b.exp(), S)
        # intra-chunk: A[t,s] = sum_d q_t k_s exp(b_t-b_s), s<=t
        diff = b.unsqueeze(1) - b.unsqueeze(0)                  # [t,s,H,D]
        C = e - s
        mask = torch.tril(torch.ones(C, C, dtype=torch.bool))
        w = torch.where(mask[:, :, None, None], diff, torch.full_like(diff, -float("inf"))).exp()
        A = torch.einsum("thd,shd,tshd->tsh", qc, kc, w)
        o_c = o_c + torch.einsum("tsh,she->the", A, vc)
        o[s:e] = o_c
        bl = b[-1]
        S = S * bl.exp().unsqueeze(-1) + torch.einsum("shd,she->hde", kc * (bl.unsqueeze(0) - b).exp(), vc)
    return o.float(), S.float()


# --------------------------------------------------------------------------- a2
def _bf16(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16)


def gla_decode_step(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, g: torch.Tensor,
                    kv_caches: torch.Tensor, slot_idx: torch.Tensor,
                    scale: Optional[float] = None, faithful: bool = True) -> torch.Tensor:
    """One decode step for B requests, state updated IN PLACE.

    Restates ``my_fused_recurrent_fwd_kernel`` (my_fused_recurrent.py:26-80):
      * slot == -1 -> request skipped (:31-35); its output row is zero here
        (the reference leaves ``torch.empty`` garbage),
      * q*scale in fp32 (:59, ``scale`` is a runtime fp32 scalar),
      * ``k (outer) v`` formed in the input dtype (bf16) BEFORE the fp32 add (:65),
      * ratio = exp(g) in fp32, new state fp32 (:69-72), output reduces the
        UN-rounded new state (:75-76), state and output stored in bf16 (:79-80).
    ``faithful=False`` gives the plain fp32 definition (no bf16 product rounding).

    q: [B,H,D]; k,v,g: [B,Hk,D]; kv_caches: [slots,Hs,D,D]; slot_idx: [B] int.
    Returns o [B,H,D] in q.dtype.
    """
    B, H, D = q.shape
    Hs = kv_caches.shape[1]
    if scale is None:
        scale = D ** -0.5
    o = torch.zeros(B, H, D, dtype=torch.float32)
    for b in range(B):
        s = int(slot_idx[b])
        if s == -1:
            continue
        kb = _expand_heads(k[b:b + 1], Hs)[0]
        vb = _expand_heads(v[b:b + 1], Hs)[0]
        gb = _expand_heads(g[b:b + 1], Hs)[0].float()
        if faithful and k.dtype == torch.bfloat16:
            kv = (kb.unsqueeze(-1) * vb.unsqueeze(-2)).float()          # bf16 product
        else:
            kv = kb.float().unsqueeze(-1) * vb.float().unsqueeze(-2)
        new = kv + gb.exp().unsqueeze(-1) * kv_caches[s].float()          # [Hs,D,D] fp32
        qb = q[b].float() * scale                                         # [H,D]
        newH = _expand_heads(new.unsqueeze(0), H)[0]
        o[b] = (qb.unsqueeze(-1) * newH).sum(-2)
        kv_caches[s] = new.to(kv_caches.dtype)
    return o.to(q.dtype)


# --------------------------------------------------------------------------- a3
def gla_prefill_varlen(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, g: torch.Tensor,
                       kv_caches: torch.Tensor, cu_seqlens: Sequence[int],
                       slot_idx: Sequence[int], init_mode: Sequence[int],
                       meta_state: Optional[torch.Tensor] = None,
                       scale: Optional[float] = None, write_state: bool = True,
                       out_dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """Prefill of N packed requests with slot states.

    Restates the request loop of ``_prefill_and_mix_infer``
    (modeling_hymeta.py:470-529; 7B twin gla_attention.py:182-211): per request
    slice ``[cu[i], cu[i+1])``, pick the initial state (zero / slot / shared
    meta state), run the recurrence (the reference's ``fused_chunk_gla``; fla
    returns the final state in fp32), store the final state into the slot in the
    cache dtype (:528), emit ``o`` token-major.
    """
    T, H, D = q.shape
    o = torch.zeros(T, H, v.shape[-1], dtype=torch.float32)
    for i in range(len(cu_seqlens) - 1):
        s, e = int(cu_seqlens[i]), int(cu_seqlens[i + 1])
        slot = int(slot_idx[i])
        mode = int(init_mode[i])
        if mode == INIT_SLOT:
            S0 = kv_caches[slot]
        elif mode == INIT_META:
            S0 = meta_state
        else:
            S0 = None
        if e > s:
            oi, Sf = gla_recurrent(q[s:e], k[s:e], v[s:e], g[s:e], S0, scale)
            o[s:e] = oi
        else:
            Hs = kv_caches.shape[1]
            Sf = torch.zeros(Hs, D, D) if S0 is None else S0.float()
        if write_state and slot >= 0:
            Hs = kv_caches.shape[1]
            if Sf.shape[0] != Hs:                     # compact (per kv-head) state
                Sf = Sf[:: Sf.shape[0] // Hs]
            kv_caches[slot] = Sf.to(kv_caches.dtype)
    return o.to(out_dtype or q.dtype)


def rms_norm_head(x: torch.Tensor, weight: torch.Tensor, eps: float) -> torch.Tensor:
    """Per-head RMSNorm applied by the 7B model after the kernel
    (gla_attention.py:210,233; hf_7B_model/norms.py:12-18)."""
    dt = x.dtype
    xf = x.float()
    var = xf.pow(2).mean(-1, keepdim=True)
    return weight * (xf * torch.rsqrt(var + eps)).to(dt)


def rms_norm_tp(x: torch.Tensor, weight: torch.Tensor, eps: float, tp_world: int = 1,
                all_reduce=None) -> torch.Tensor:
    """``HymetaRMSNormTP._forward`` (vllm_hymeta/models/modeling_hymeta.py:151-173) restated:
    fp32 x, variance = mean over the LOCAL hidden slice, summed over the tensor-parallel ranks by
    ``all_reduce`` and divided by tp_world, x * rsqrt(var + eps), cast back to the input dtype, times
    the rank's fp32 weight slice (result promoted to fp32).  Pinned bit-exactly to the unmodified
    class at TP world 1 / 2 / 4 (tests/golden/hymeta_rmsnorm_tp.npz, made by oracle/make_golden.py
    with the vLLM names the module imports stubbed; tests/test_oracle_golden.py)."""
    orig = x.dtype
    xf = x.to(torch.float32)
    var = xf.pow(2).mean(dim=-1, keepdim=True, dtype=torch.float32)
    if tp_world > 1:
        var = all_reduce(var) / tp_world
    xf = xf * torch.rsqrt(var + eps)
    return xf.to(orig) * weight[: x.shape[-1]]


def branch_norm_merge(out_attn, out_linear, w1, w2, eps, tp_world: int = 1, all_reduce=(None, None)):
    """modeling_hymeta.py:909-919: (norm1(out_attn) + norm2(out_linear)) / 2 -> bfloat16."""
    a = rms_norm_tp(out_attn, w1, eps, tp_world, all_reduce[0])
    b = rms_norm_tp(out_linear, w2, eps, tp_world, all_reduce[1])
    return ((a + b) / 2).to(torch.bfloat16)


# ---------------------------------------------------------------- synthetic inputs
def synth_gla_inputs(T: int, H: int, Hkv: int, D: int, model: str, seed: int,
                     dtype=torch.bfloat16):
    """SURVEY 8(d) synthetic inputs: q,k,v ~ N(0,1) then the reference's own
    pre-processing (a4).  Returns q [T,H,D], k,v [T,Hkv,D], g [T,Hkv,D]
    (fp32 for 70B, ``dtype`` for 7B)."""
    gen = torch.Generator().manual_seed(seed)
    q = torch.randn(T, H, D, generator=gen).to(dtype)
    k = torch.randn(T, Hkv, D, generator=gen).to(dtype)
    v = torch.randn(T, Hkv, D, generator=gen).to(dtype)
    if model == "7b":
        gk = torch.randn(T, Hkv, D, generator=gen).to(dtype)
        q, k, g = gate_7b(q, k, gk)
    elif model == "70b":
        q, k, g = gate_70b(q, k, v)
    else:
        raise ValueError(model)
    return q, k, v, g
