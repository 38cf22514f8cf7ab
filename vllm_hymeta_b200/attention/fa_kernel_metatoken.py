"""Kernel-level functions of the meta-token attention path -- same names, argument meaning and
assertions as the reference vllm_hymeta/attention/fa_kernel_metatoken.py:287-374, routed to the
single-pass sm_100a kernels (no flash-attn, no LSE-merge launches)."""
from __future__ import annotations

from typing import Optional

import torch

from .. import ops


def _window_left(window_size) -> int:
    """flash-attn convention: window_size[0] = keys strictly before the query that stay
    visible; -1 = unbounded (reference meta_attention.py:671-672)."""
    return int(window_size[0]) if window_size is not None else -1


def _check_unused(dropout_p, alibi_slopes, softcap, attn_mask, return_attn_probs):
    # accepted for signature compatibility; every reference configuration passes the defaults
    # (meta_attention.py:866-884: dropout 0.0, alibi None, softcap 0)
    if dropout_p not in (0, 0.0):
        raise NotImplementedError("dropout is not supported (inference path)")
    if alibi_slopes is not None:
        raise NotImplementedError("alibi_slopes is not supported by the Hymeta kernels")
    if softcap not in (0, 0.0, None):
        raise NotImplementedError("logits soft cap is not supported by the Hymeta kernels")
    if attn_mask is not None or return_attn_probs:
        raise NotImplementedError("attn_mask / return_attn_probs are not supported")


def metatoken_flash_attn_varlen_func(
    q1, q2, k1, k2, v1, v2,
    cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=None, max_seqlen_k=None,
    num_meta_tokens=128, dropout_p=0.0, softmax_scale=None, causal=False,
    window_size=(-1, -1), alibi_slopes=None, block_tables=None, softcap=0.0,
    deterministic=False, return_attn_probs=False, attn_mask=None, out: Optional[torch.Tensor] = None,
):
    """Packed varlen prefill (reference fa_kernel_metatoken.py:329-374 -> :198-285).
    Returns ``[1, (M if q2) + Tq, H, D]`` with the meta-query rows FIRST (:283).  Extension over the
    reference signature: ``out`` [(M) + Tq, H, D] is written in place (the backend passes its output
    tensor, which saves the reference's full copy ``output[:n] = res[0]``, meta_attention.py:920)."""
    _check_unused(dropout_p, alibi_slopes, softcap, attn_mask, return_attn_probs)
    assert causal and window_size[1] in [0, -1]
    if k2 is not None:
        assert num_meta_tokens == k2.shape[0]
    if q2 is not None:
        assert q1.shape[0] == k1.shape[0] or block_tables is not None, \
            "meta_tokens' query doesn't support decoding"
        assert num_meta_tokens == q2.shape[0]
    if softmax_scale is None:
        softmax_scale = q1.shape[-1] ** (-0.5)
    if max_seqlen_q is None:
        max_seqlen_q = q1.shape[0]
    out = ops.swa_meta_prefill_varlen(
        q1, k1, v1, q2, k2, v2, cu_seqlens_q, cu_seqlens_k, int(max_seqlen_q),
        _window_left(window_size), scale=softmax_scale, block_tables=block_tables, out=out)
    return out.unsqueeze(0)


def metatoken_flash_attn_with_kvcache(
    q1, q2, k2, v2, key_cache=None, value_cache=None, cache_seqlens=None,
    num_meta_tokens=128, dropout_p=0.0, softmax_scale=None, causal=False,
    window_size=(-1, -1), alibi_slopes=None, block_tables=None, softcap=0.0,
    deterministic=False, return_attn_probs=False, attn_mask=None, out: Optional[torch.Tensor] = None,
    new_key: Optional[torch.Tensor] = None, new_value: Optional[torch.Tensor] = None,
    new_slot_mapping: Optional[torch.Tensor] = None,
):
    """Single-token paged decode (reference fa_kernel_metatoken.py:287-326 -> :124-194).
    q1 [B,H,D] -> ``[B, 1, H, D]``.  Extension over the reference signature: ``new_key`` /
    ``new_value`` [B,Hkv,D] + ``new_slot_mapping`` [B] fold the step's ``reshape_and_cache_flash``
    (meta_attention.py:790-799) into the kernel."""
    _check_unused(dropout_p, alibi_slopes, softcap, attn_mask, return_attn_probs)
    assert causal and window_size[1] in [0, -1]
    if k2 is not None:
        assert num_meta_tokens == k2.shape[0]
    assert q2 is None, "meta_tokens' query doesn't support decoding"
    if softmax_scale is None:
        softmax_scale = q1.shape[-1] ** (-0.5)
    o = ops.swa_meta_decode_paged(q1, key_cache, value_cache, block_tables, cache_seqlens, k2, v2,
                                  _window_left(window_size), scale=softmax_scale, out=out,
                                  new_k=new_key, new_v=new_value, new_slot_mapping=new_slot_mapping)
    return o.unsqueeze(1)
