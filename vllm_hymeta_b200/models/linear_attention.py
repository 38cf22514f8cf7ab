"""GLA attention core of the 70B model -- the part of the reference ``HLinearAttention``
(vllm_hymeta/models/modeling_hymeta.py:366-659) between ``qkv_proj`` and the layer's norm, with the
same method names and argument meaning, on the sm_100a kernels:

* gate / feature map (:598-629)            -> ONE launch, ``ops.gla_gate_70b`` (no repeat_kv copies)
* meta pass (:447-467)                     -> ``ops.gla_prefill_varlen`` on the meta tokens, state
                                              written straight into ``meta_linear_cache``
* per-request ``fused_chunk_gla`` loop     -> ONE varlen launch over every prefill request; the
  (:470-529)                                  initial state (slot / shared meta state) is chosen on
                                              the device from ``context_lens_tensor`` -- the two
                                              host syncs per request per layer (:489,501) are gone
* ``my_fused_recurrent_gla`` decode (:551-578) -> ``ops.gla_decode`` on views (no ``.contiguous()``)

Projections (``QKVParallelLinear``) are vLLM-side and out of scope: ``forward``
takes their output.  Token layout of one step, as the reference builds it (:441-549):
``[meta tokens (only while the meta caches warm up) | prefill tokens | decode tokens]``.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import ops
from .hymeta_cache import HymetaCacheParams


class HLinearAttention:
    """Rank-local GLA core: ``num_heads`` / ``num_key_value_heads`` are the TOTAL counts and are
    divided by ``tp_size`` exactly as the reference does (:401-418)."""

    def __init__(self, hidden_size: int = 1024, num_heads: Optional[int] = None,
                 num_key_value_heads: Optional[int] = None, num_meta_tokens: int = 0,
                 clamp_max: Optional[float] = 0.95, layer_idx: Optional[int] = None,
                 tp_size: int = 1, tp_rank: int = 0):
        self.hidden_size = hidden_size
        self.num_heads = num_heads
        self.total_num_key_value_heads = num_key_value_heads
        self.num_key_value_groups = num_heads // num_key_value_heads
        self.num_meta_tokens = num_meta_tokens
        self.key_dim = hidden_size
        self.head_dim = hidden_size // num_heads
        self.clamp_max = clamp_max
        self.layer_idx = layer_idx
        assert self.key_dim % num_heads == 0, f"key dim must be divisible by num_heads of {num_heads}"
        self.tp_size, self.tp_rank = tp_size, tp_rank
        assert num_heads % tp_size == 0, "num_heads must be divisible by tp_size"
        self.tp_heads = num_heads // tp_size
        if num_key_value_heads >= tp_size:
            assert num_key_value_heads % tp_size == 0
        else:
            assert tp_size % num_key_value_heads == 0
        self.tp_kv_heads = max(1, num_key_value_heads // tp_size)
        self.q_size = self.tp_heads * self.head_dim
        self.kv_size = self.tp_kv_heads * self.head_dim

    # ------------------------------------------------------------------------------ a4
    def _gate(self, qkv: torch.Tensor, lower_bound: Optional[torch.Tensor]):
        """q, k, v views of the projection output -> (relu(q), k', v, g) with k'/g per kv-head
        (no lower bound) or per q-head (lower bound), as at :598-629."""
        T = qkv.shape[0]
        q = qkv[:, :self.q_size].view(T, self.tp_heads, self.head_dim)
        k = qkv[:, self.q_size:self.q_size + self.kv_size].view(T, self.tp_kv_heads, self.head_dim)
        v = qkv[:, self.q_size + self.kv_size:].view(T, self.tp_kv_heads, self.head_dim)
        lb = None
        if lower_bound is not None and self.layer_idx % 7 != 0:
            # rank-local heads of the per-q-head row (:620-625)
            lb = lower_bound.reshape(self.num_heads, self.head_dim)
            lb = lb[self.tp_rank * self.tp_heads:(self.tp_rank + 1) * self.tp_heads]
            lb = lb.to(torch.float32).contiguous()
        q, k, g = ops.gla_gate_70b(q, k, lb, clamp_max=self.clamp_max)
        return q, k, v, g

    # ------------------------------------------------------------------------------ a3
    def _prefill_and_mix_infer(self, q, k, v, g, kv_cache, has_meta_cache, meta_cache,
                               state_indices_tensor, attn_metadata, is_vllm_testing):
        """q [T,H,D]; k,g [T,Hk,D]; v [T,Hv,D]; kv_cache [slots,Hs,D,D]; meta_cache [Hs,D,D].
        Returns [T, H*D]."""
        T, H, D = q.shape
        out = torch.empty(T, H, D, dtype=q.dtype, device=q.device)
        num_prefills = getattr(attn_metadata, "num_prefills", 0)
        warm = (not has_meta_cache) and (not is_vllm_testing)
        off = self.num_meta_tokens if warm else 0
        i32 = dict(dtype=torch.int32, device=q.device)
        if warm and off > 0:
            # meta pass: zero initial state, final state IS the shared meta state (:447-467)
            ops.gla_prefill_varlen(
                q[:off], k[:off], v[:off], g[:off], meta_cache.unsqueeze(0),
                torch.tensor([0, off], **i32), torch.zeros(1, **i32), torch.zeros(1, **i32),
                write_state=True, max_seqlen=off, out=out[:off])
        if num_prefills > 0:
            n_tok = attn_metadata.num_prefill_tokens
            cu = attn_metadata.query_start_loc[:num_prefills + 1].to(torch.int32)
            slots = state_indices_tensor[:num_prefills].to(torch.int32)
            if is_vllm_testing:
                mode = torch.full((num_prefills,), ops.INIT_SLOT, **i32)
            else:
                # context_len == 0 -> first chunk of the request: start from the meta state (:501-504)
                ctx = attn_metadata.context_lens_tensor[:num_prefills]
                mode = torch.where(ctx == 0, ops.INIT_META, ops.INIT_SLOT).to(torch.int32)
            sl = slice(off, off + n_tok)
            ops.gla_prefill_varlen(
                q[sl], k[sl], v[sl], g[sl], kv_cache, cu, slots, mode, meta_state=meta_cache,
                write_state=not is_vllm_testing,
                max_seqlen=getattr(attn_metadata, "max_query_len", None) or n_tok, out=out[sl])
        if attn_metadata.num_decode_tokens > 0:
            start = attn_metadata.num_prefill_tokens + off
            self._decode(q, k, v, g, kv_cache, has_meta_cache, state_indices_tensor, attn_metadata,
                         is_vllm_testing, out=out[start:])
        return out.view(T, H * D)

    # ------------------------------------------------------------------------------ a2
    def _decode(self, q, k, v, g, kv_cache, has_meta_cache, state_indices_tensor, attn_metadata,
                is_vllm_testing, out=None):
        off = self.num_meta_tokens if (not has_meta_cache and not is_vllm_testing) else 0
        start = attn_metadata.num_prefill_tokens + off
        slot_ids = state_indices_tensor[getattr(attn_metadata, "num_prefills", 0):]
        B = q.shape[0] - start
        o = ops.gla_decode(q[start:], k[start:], v[start:], g[start:], kv_cache,
                           slot_ids[:B].to(torch.int32), out=out)
        return o.view(B, -1)

    def forward(self, qkv: torch.Tensor, kv_caches: HymetaCacheParams, attn_metadata,
                has_meta_cache: Optional[bool] = False, lower_bound: Optional[torch.Tensor] = None,
                is_vllm_testing: Optional[bool] = False) -> torch.Tensor:
        """``qkv``: output of the qkv projection, [num_tokens, (H + 2 Hkv) * D] rank-local."""
        q, k, v, g = self._gate(qkv, lower_bound)
        kv_cache = kv_caches.hymeta_cache
        meta_cache = kv_caches.meta_linear_cache
        state_indices_tensor = kv_caches.state_indices_tensor
        decode_only = getattr(attn_metadata, "num_prefills", 0) == 0
        if not decode_only or not has_meta_cache:
            return self._prefill_and_mix_infer(q, k, v, g, kv_cache, has_meta_cache, meta_cache,
                                               state_indices_tensor, attn_metadata, is_vllm_testing)
        return self._decode(q, k, v, g, kv_cache, has_meta_cache, state_indices_tensor,
                            attn_metadata, is_vllm_testing)
