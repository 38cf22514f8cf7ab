"""GLA recurrent decode -- same name and arguments as the reference
vllm_hymeta/models/my_fused_recurrent_gla.py:8-23 (-> my_fused_recurrent.py:83-135)."""
from __future__ import annotations

import torch

from .. import ops


def _decode(q, k, v, g, kv_caches, slot_idx, scale):
    B, H, L, D = q.shape
    # used for decoding, so the sequence length is 1 (reference my_fused_recurrent.py:92-93)
    assert L == 1 and k.shape[2] == 1 and v.shape[2] == 1
    assert k.shape == v.shape
    if slot_idx.dtype != torch.int32:
        slot_idx = slot_idx.to(torch.int32)
    return ops.gla_decode(q[:, :, 0], k[:, :, 0], v[:, :, 0], g[:, :, 0], kv_caches, slot_idx, scale)


def my_fused_recurrent_gla(q, k, v, g=None, kv_caches=None, slot_idx=None, scale=None,
                           initial_state=None, output_final_state=False, reverse=False):
    """q [B,H,1,D]; k,v,g [B,Hk,1,D] (Hk = H as the reference passes them after repeat_kv, or
    the un-repeated kv-heads); kv_caches [slots,Hs,D,D] updated in place; slot_idx [B] int32,
    -1 = skip.  Returns ``[B, H*D]`` (70B variant, my_fused_recurrent.py:117-119)."""
    if scale is None:
        scale = q.shape[-1] ** -0.5
    o = _decode(q, k, v, g, kv_caches, slot_idx, scale)
    return o.reshape(o.shape[0], -1)
