"""7B twin of the GLA recurrent decode (reference vllm_hymeta/model_for_7B/my_fused_recurrent.py:
83-153): identical kernel, returns ``[B, H, D]`` so the per-head g_norm can follow."""
from __future__ import annotations

from ..models.my_fused_recurrent_gla import _decode


def my_fused_recurrent_gla(q, k, v, g=None, kv_caches=None, slot_idx=None, scale=None,
                           initial_state=None, output_final_state=False, reverse=False):
    if scale is None:
        scale = q.shape[-1] ** -0.5
    return _decode(q, k, v, g, kv_caches, slot_idx, scale)


my_fused_recurrent = my_fused_recurrent_gla
