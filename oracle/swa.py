"""Oracle: sliding-window attention with the meta-token prefix.
TEST INFRASTRUCTURE -- see oracle/__init__.py.

Definition (vllm_hymeta/attention/fa_kernel_metatoken.py:377-410
``naive_metatoken_flash_attn``): a real query at position ``i`` attends to
``[meta keys 0..M-1 | real keys j <= i with i - j <= window_left]`` with one
joint softmax, scale ``D**-0.5``; meta keys are always visible.  Meta queries
(present only on the pass that fills the shared caches) attend to the meta keys
only, bidirectionally (:396-400, ``causal=False`` at :180,:267).

``window_left`` = number of keys strictly before the query that stay visible
(flash-attn's ``window_size_left``): 70B passes ``W-1`` (meta_attention.py:671-672),
7B passes ``W`` (hf_7B_model/window_attention.py:215), full-attention layers -1.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F


def _rep(x: torch.Tensor, H: int) -> torch.Tensor:
    """[L, Hkv, D] -> [L, H, D] (GQA expansion, interleaved like repeat_kv)."""
    return x if x.shape[1] == H else x.repeat_interleave(H // x.shape[1], dim=1)


def _attend(q, k, v, mask, scale):
    """q [Lq,H,D], k,v [Lk,H,D], mask [Lq,Lk] bool -> (out [Lq,H,D] fp32, lse [H,Lq])."""
    s = torch.einsum("qhd,khd->hqk", q.float(), k.float()) * scale
    s = torch.where(mask[None], s, torch.full_like(s, -float("inf")))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)          # fully masked row -> zeros (flash-attn behaviour)
    return torch.einsum("hqk,khd->qhd", p, v.float()), lse


def window_mask(Lq: int, Lk: int, window_left: int) -> torch.Tensor:
    """Causal + left-window mask with queries aligned to the END of the keys
    (fa_kernel_metatoken.py:387-392; flash-attn bottom-right alignment)."""
    q_id = Lk - Lq + torch.arange(Lq)
    k_id = torch.arange(Lk)
    m = q_id[:, None] >= k_id[None, :]
    if window_left >= 0:
        m &= (q_id[:, None] - k_id[None, :]) <= window_left
    return m


def swa_meta_dense(q1: torch.Tensor, k1: torch.Tensor, v1: torch.Tensor,
                   q2: Optional[torch.Tensor], k2: Optional[torch.Tensor],
                   v2: Optional[torch.Tensor], window_left: int,
                   scale: Optional[float] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """ONE sequence.  q1 [Lq,H,D], k1,v1 [Lk,Hkv,D], k2,v2 [M,Hkv,D] or None,
    q2 [M,H,D] or None.  Returns (out_real [Lq,H,D] fp32, out_meta [M,H,D] fp32|None)."""
    Lq, H, D = q1.shape
    Lk = k1.shape[0]
    if scale is None:
        scale = D ** -0.5
    M = 0 if k2 is None else k2.shape[0]
    mask = window_mask(Lq, Lk, window_left)
    k, v = _rep(k1, H), _rep(v1, H)
    if M:
        mask = torch.cat([torch.ones(Lq, M, dtype=torch.bool), mask], dim=1)
        k = torch.cat([_rep(k2, H), k], 0)
        v = torch.cat([_rep(v2, H), v], 0)
    out, _ = _attend(q1, k, v, mask, scale)
    out_meta = None
    if q2 is not None:
        out_meta, _ = _attend(q2, _rep(k2, H), _rep(v2, H),
                              torch.ones(q2.shape[0], M, dtype=torch.bool), scale)
    return out, out_meta


def merge_out_lse(out, lse, block_out, block_lse):
    """``_update_out_and_lse`` (fa_kernel_metatoken.py:110-122) restated.
    out [B,L,H,D], lse [B,H,L]."""
    o, bo = out.float(), block_out.float()
    new_lse = lse - F.logsigmoid(lse - block_lse)
    o = o - torch.sigmoid(block_lse - lse).transpose(1, 2)[..., None] * (o - bo)
    return o.to(out.dtype), new_lse


def swa_meta_two_pass(q1, k1, v1, k2, v2, window_left, scale=None):
    """The reference's actual evaluation order for one sequence: windowed causal
    pass + non-causal meta pass + LSE merge (fa_kernel_metatoken.py:235-277).
    Must equal swa_meta_dense; used to pin the merge algebra."""
    Lq, H, D = q1.shape
    if scale is None:
        scale = D ** -0.5
    o1, l1 = _attend(q1, _rep(k1, H), _rep(v1, H), window_mask(Lq, k1.shape[0], window_left), scale)
    o2, l2 = _attend(q1, _rep(k2, H), _rep(v2, H),
                     torch.ones(Lq, k2.shape[0], dtype=torch.bool), scale)
    o, _ = merge_out_lse(o1[None], l1[None], o2[None], l2[None])
    return o[0]


# --------------------------------------------------------------------------- a7
def swa_meta_prefill_varlen(q1, k1, v1, q2, k2, v2, cu_seqlens_q: Sequence[int],
                            cu_seqlens_k: Sequence[int], window_left: int,
                            scale: Optional[float] = None, out_dtype=None) -> torch.Tensor:
    """Packed varlen prompts (``FlashAttentionVarlenWithMetaToken.forward``
    fa_kernel_metatoken.py:198-285).  Sequence i has queries
    q1[cu_q[i]:cu_q[i+1]] and keys k1[cu_k[i]:cu_k[i+1]] (queries aligned to the
    end of the keys).  Output rows: META ROWS FIRST when q2 is given (:283)."""
    outs = []
    for i in range(len(cu_seqlens_q) - 1):
        qs, qe = int(cu_seqlens_q[i]), int(cu_seqlens_q[i + 1])
        ks, ke = int(cu_seqlens_k[i]), int(cu_seqlens_k[i + 1])
        o, _ = swa_meta_dense(q1[qs:qe], k1[ks:ke], v1[ks:ke], None, k2, v2, window_left, scale)
        outs.append(o)
    out = torch.cat(outs, 0) if outs else torch.zeros(0, *q1.shape[1:])
    if q2 is not None:
        H = q2.shape[1]
        om, _ = _attend(q2, _rep(k2, H), _rep(v2, H),
                        torch.ones(q2.shape[0], k2.shape[0], dtype=torch.bool),
                        scale if scale is not None else q2.shape[-1] ** -0.5)
        out = torch.cat([om, out], 0)
    return out.to(out_dtype or q1.dtype)


# --------------------------------------------------------------------------- a9
def kv_write_paged(key, value, k_cache, v_cache, slot_mapping) -> None:
    """``reshape_and_cache_flash`` (meta_attention.py:790-799) restated:
    cache[slot // bs, slot % bs, h, :] = key[t, h, :] for slot >= 0.  In place."""
    bs = k_cache.shape[1]
    for t in range(key.shape[0]):
        s = int(slot_mapping[t])
        if s < 0:
            continue
        k_cache[s // bs, s % bs] = key[t].to(k_cache.dtype)
        v_cache[s // bs, s % bs] = value[t].to(v_cache.dtype)


def gather_paged(cache, block_table_row, lo: int, hi: int) -> torch.Tensor:
    """Logical positions [lo, hi) of one sequence from cache [nb, bs, Hkv, D]."""
    bs = cache.shape[1]
    pos = torch.arange(lo, hi)
    blk = torch.as_tensor(block_table_row, dtype=torch.long)[pos // bs]
    return cache[blk, pos % bs]


# --------------------------------------------------------------------------- a8
def swa_meta_decode_paged(q, k_cache, v_cache, block_tables, cache_seqlens, k2, v2,
                          window_left: int, scale: Optional[float] = None,
                          out_dtype=None) -> torch.Tensor:
    """One new token per request over the paged cache
    (``FlashAttentionWithMetaToken.forward`` fa_kernel_metatoken.py:124-194:
    ``fwd_kvcache`` with window ``(window_left, -1)``, causal, seqlen_q = 1, plus
    the meta pass and the merge).  ``cache_seqlens[b]`` counts the keys already
    in the cache INCLUDING the current token (the KV write precedes attention,
    meta_attention.py:790-799).  q [B,H,D] -> out [B,H,D]."""
    B, H, D = q.shape
    out = torch.zeros(B, H, D, dtype=torch.float32)
    for b in range(B):
        L = int(cache_seqlens[b])
        lo = 0 if window_left < 0 else max(0, L - 1 - window_left)
        kk = gather_paged(k_cache, block_tables[b], lo, L)
        vv = gather_paged(v_cache, block_tables[b], lo, L)
        if L == 0 and (k2 is None or k2.shape[0] == 0):
            continue
        o, _ = swa_meta_dense(q[b:b + 1], kk, vv, None, k2, v2, -1, scale)
        out[b] = o[0]
    return out.to(out_dtype or q.dtype)


def swa_meta_prefill_paged(q1, k_cache, v_cache, block_tables, query_start_loc,
                           seq_lens, k2, v2, window_left, scale=None, out_dtype=None):
    """Prefix-cache / chunked-prefill branch (meta_attention.py:886-919), evident
    intent: request i has ``seq_lens[i]`` keys in the paged cache and
    its ``query_start_loc[i+1]-query_start_loc[i]`` queries are the LAST tokens."""
    outs = []
    for i in range(len(seq_lens)):
        qs, qe = int(query_start_loc[i]), int(query_start_loc[i + 1])
        L = int(seq_lens[i])
        kk = gather_paged(k_cache, block_tables[i], 0, L)
        vv = gather_paged(v_cache, block_tables[i], 0, L)
        o, _ = swa_meta_dense(q1[qs:qe], kk, vv, None, k2, v2, window_left, scale)
        outs.append(o)
    return torch.cat(outs, 0).to(out_dtype or q1.dtype)


# ------------------------------------------------------------------------ RoPE
def rope_neox(x: torch.Tensor, positions: torch.Tensor, theta: float = 1e6) -> torch.Tensor:
    """NeoX rotary embedding in fp32 (hf_model/attention.py:26-82 /
    hf_7B_model/window_attention.py:23-79).  x [T,H,D], positions [T]."""
    D = x.shape[-1]
    inv = 1.0 / (theta ** (torch.arange(0, D, 2, dtype=torch.int64).float() / D))
    fr = positions.float()[:, None] * inv[None]
    emb = torch.cat([fr, fr], -1)
    cos, sin = emb.cos().to(x.dtype)[:, None], emb.sin().to(x.dtype)[:, None]
    x1, x2 = x[..., : D // 2], x[..., D // 2:]
    return x * cos + torch.cat([-x2, x1], -1) * sin


def rope_cos_sin_cache(D: int, max_pos: int, base: float, dtype=torch.bfloat16) -> torch.Tensor:
    """HymetaRotaryEmbedding._compute_cos_sin_cache (vllm_hymeta/models/modeling_hymeta.py:205-218):
    [max_pos, D] = [cos | sin] of t * inv_freq, fp32 then cast to cache_dtype (bf16, :1187)."""
    inv_freq = 1.0 / (base ** (torch.arange(0, D, 2, dtype=torch.float) / D))
    freqs = torch.einsum("i,j -> ij", torch.arange(max_pos, dtype=torch.float), inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


def rope_neox_table(x: torch.Tensor, positions: torch.Tensor, cos_sin: torch.Tensor) -> torch.Tensor:
    """vLLM's rotary_embedding op (vllm==0.10.0 csrc/pos_encoding_kernels.cu, NeoX branch; call
    site modeling_hymeta.py:230) in the dtype of ``x``: x' = x*cos - y*sin, y' = y*cos + x*sin with
    x = head[:D/2], y = head[D/2:], every product and sum rounded to x.dtype (c10::BFloat16
    operators round after each operation; torch CPU bf16 arithmetic does the same).  x [T,H,D]."""
    D = x.shape[-1]
    cs = cos_sin[positions.long()].to(x.dtype)
    cos, sin = cs[:, None, : D // 2], cs[:, None, D // 2:]
    a, b = x[..., : D // 2], x[..., D // 2:]
    return torch.cat([a * cos - b * sin, b * cos + a * sin], -1)


def rope_kv_prologue(q, k, v, cos_sin, positions, pos_offset, k_cache, v_cache, slot_mapping,
                     meta_k=None, meta_v=None, n_meta=0):
    """The SWA layer prologue around the attention call (modeling_hymeta.py:1314-1324 positions,
    :762 rotary, :776-792 meta capture, meta_attention.py:790-799 page write), restated step by
    step.  Returns rotated (q, k); caches / meta_k / meta_v are updated in place."""
    T = q.shape[0]
    pos = torch.cat([torch.arange(n_meta, dtype=torch.int64), positions[: T - n_meta] + pos_offset])
    q, k = rope_neox_table(q, pos, cos_sin), rope_neox_table(k, pos, cos_sin)
    if n_meta:
        meta_k[:n_meta].copy_(k[:n_meta])
        meta_v[:n_meta].copy_(v[:n_meta])
    if k_cache is not None:
        kv_write_paged(k[n_meta:], v[n_meta:], k_cache, v_cache, slot_mapping)
    return q, k


def synth_swa_inputs(T: int, H: int, Hkv: int, D: int, seed: int, M: int = 0,
                     pos_offset: int = 0, dtype=torch.bfloat16, theta: float = 1e6):
    """q,k,v ~ N(0,1), RoPE with positions M + t; meta K/V ~ N(0,1)."""
    gen = torch.Generator().manual_seed(seed)
    q = torch.randn(T, H, D, generator=gen).to(dtype)
    k = torch.randn(T, Hkv, D, generator=gen).to(dtype)
    v = torch.randn(T, Hkv, D, generator=gen).to(dtype)
    pos = torch.arange(T) + M + pos_offset
    q, k = rope_neox(q, pos, theta), rope_neox(k, pos, theta)
    k2 = v2 = None
    if M:
        k2 = torch.randn(M, Hkv, D, generator=gen).to(dtype)
        v2 = torch.randn(M, Hkv, D, generator=gen).to(dtype)
    return q, k, v, k2, v2


# ---------------------------------------------------------------------------------- a10
def advance_step(num_seqs, num_queries, block_size, input_tokens, sampled_token_ids,
                 input_positions, seq_lens, slot_mapping, block_tables) -> None:
    """vLLM 0.10.0 ``advance_step_flashattn`` (csrc/prepare_inputs/advance_step.cu; not vendored in
    the reference, call site vllm_hymeta/attention/meta_attention.py:401-409) restated, in place:
    live rows take their sampled token, grow by one position and get that position's slot through
    the block table; CUDA-graph padding rows get slot -1."""
    for i in range(num_seqs):
        if i >= num_queries:
            slot_mapping[i] = -1
            continue
        if input_tokens is not None and sampled_token_ids is not None:
            input_tokens[i] = sampled_token_ids[i]
        seq_lens[i] += 1
        pos = int(seq_lens[i]) - 1
        if input_positions is not None:
            input_positions[i] = pos
        slot_mapping[i] = int(block_tables[i][pos // block_size]) * block_size + pos % block_size
