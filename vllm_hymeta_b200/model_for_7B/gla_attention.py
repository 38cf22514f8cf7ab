"""GLA attention core of the 7B model -- the part of the reference ``GatedLinearAttention``
(vllm_hymeta/model_for_7B/gla_attention.py:90-283) between the projections and ``o_proj``, same
method names and argument meaning, on the sm_100a kernels:

* feature map + gate (:251-267)          -> folded into the GLA kernels: they read the
                                            raw ``qkv_proj`` / ``gk_proj`` rows, relu(q) / relu(k)
                                            on the fly; the decode kernel computes logsigmoid(gk) / 16
                                            itself, the prefill kernel's gate warps look it up in a
                                            64 K-entry table (the gate is a pure function of a bf16
                                            input); k / v / gk stay at the kv-head count: no
                                            ``repeat_kv`` x7 materialisation
* per-request ``fused_chunk_gla`` loop   -> ONE varlen launch over all prefill requests, slot
  (:179-211)                                states read and written in-kernel (no isnan() host sync)
* decode (:223-235)                      -> ``ops.gla_decode``
* ``g_norm`` per-head RMSNorm (:210,233) -> fused into both kernels as their epilogue
"""
from __future__ import annotations

import math
from typing import Optional

import torch

from .. import ops
from .gla_cache import GLACacheParams


class GatedLinearAttention:

    def __init__(self, hidden_size: int, num_heads: int, num_kv_heads: int,
                 gate_logit_normalizer: float = 16.0, norm_eps: float = 1e-6, tp_size: int = 1,
                 tp_rank: int = 0, g_norm_weight: Optional[torch.Tensor] = None):
        self.hidden_size = hidden_size
        self.num_heads = num_heads // tp_size
        self.total_num_kv_heads = num_kv_heads
        self.tp_size, self.tp_rank = tp_size, tp_rank
        assert num_heads % tp_size == 0
        self.tp_kv_heads = max(1, num_kv_heads // tp_size)
        self.num_key_value_groups = self.num_heads // self.tp_kv_heads
        self.head_dim = hidden_size // num_heads
        self.q_size = self.num_heads * self.head_dim
        self.kv_size = self.tp_kv_heads * self.head_dim
        self.gate_logit_normalizer = gate_logit_normalizer
        self.norm_eps = norm_eps
        self.g_norm_weight = g_norm_weight            # [head_dim]

    def g_norm(self, x: torch.Tensor) -> torch.Tensor:
        """Per-head RMSNorm (reference hf_7B_model/norms.py:12-18)."""
        dt = x.dtype
        xf = x.float()
        xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.norm_eps)
        w = self.g_norm_weight if self.g_norm_weight is not None else 1.0
        return w * xf.to(dt)

    def _gnorm_w(self, device) -> torch.Tensor:
        """bf16 [head_dim] weight handed to the kernels' fused g_norm epilogue."""
        w = self.g_norm_weight
        if w is None:
            w = self.g_norm_weight = torch.ones(self.head_dim, dtype=torch.bfloat16, device=device)
        elif w.dtype != torch.bfloat16 or w.device != device:
            w = self.g_norm_weight = w.to(device=device, dtype=torch.bfloat16)
        return w

    def _prefill_and_mix_infer(self, q, k, v, gk, kv_cache, state_indices_tensor, attn_metadata):
        T, H, D = q.shape
        out = torch.empty(T, H, D, dtype=q.dtype, device=q.device)
        num_prefills = getattr(attn_metadata, "num_prefills", 0)
        if num_prefills > 0:
            n_tok = attn_metadata.num_prefill_tokens
            cu = attn_metadata.query_start_loc[:num_prefills + 1].to(torch.int32)
            slots = state_indices_tensor[:num_prefills].to(torch.int32)
            # context_len == 0: a new (or recompute-preempted) request starts from a ZERO state
            # whatever its possibly recycled slot holds -- the reference zeroes those slots in
            # GLAswaModel._clear_prefill_cache (modeling_gla_swa.py:391-414, called at :450); here
            # the choice is made on device per request, no host sync and no separate zero-fill
            ctx = getattr(attn_metadata, "context_lens_tensor", None)
            if ctx is not None:
                mode = torch.where(ctx[:num_prefills] == 0, ops.INIT_ZERO, ops.INIT_SLOT).to(torch.int32)
            else:
                mode = torch.full((num_prefills,), ops.INIT_SLOT, dtype=torch.int32, device=q.device)
            # ONE launch: relu(q), relu(k) on the fly, the gate looked up from the 64 K-entry table of
            # bf16(bf16(logsigmoid(x)) / normalizer) by the kernel's gate warps
            ops.gla_prefill_varlen(q[:n_tok], k[:n_tok], v[:n_tok], gk[:n_tok], kv_cache, cu, slots,
                                   mode, max_seqlen=getattr(attn_metadata, "max_query_len", None) or n_tok,
                                   out=out[:n_tok], gnorm_weight=self._gnorm_w(q.device),
                                   gnorm_eps=self.norm_eps, relu_qk=True,
                                   gate_table=ops.gla_gate_7b_table(self.gate_logit_normalizer, q.device))
        if attn_metadata.num_decode_tokens > 0:
            self._decode(q, k, v, gk, kv_cache, state_indices_tensor, attn_metadata,
                         out=out[attn_metadata.num_prefill_tokens:])
        return out.view(T, H * D)                  # g_norm already applied by both kernels' epilogues

    def _decode(self, q, k, v, gk, kv_cache, state_indices_tensor, attn_metadata, out=None,
                norm=True):
        start = attn_metadata.num_prefill_tokens
        slot_ids = state_indices_tensor[getattr(attn_metadata, "num_prefills", 0):]
        B = q.shape[0] - start
        qd, kd, gd, fused = q[start:], k[start:], gk[start:], self.gate_logit_normalizer
        if math.frexp(float(fused))[0] != 0.5:
            # the in-kernel gate divides by an exact power of two only (the model's 16): any other
            # normalizer goes through the gate kernel first
            qd, kd, gd = ops.gla_gate_7b(qd, kd, gd, fused)
            fused = None
        o = ops.gla_decode(qd, kd, v[start:], gd, kv_cache,
                           slot_ids[:B].to(torch.int32), out=out,
                           gnorm_weight=self._gnorm_w(q.device) if norm else None,
                           gnorm_eps=self.norm_eps, gate_normalizer=fused)
        return o.view(B, -1)

    def forward(self, qkv: torch.Tensor, gk: torch.Tensor, kv_caches: GLACacheParams,
                attn_metadata) -> torch.Tensor:
        """``qkv``: [num_tokens, (H + 2 Hkv) * D] rank-local projection output; ``gk``: output of
        ``gk_proj`` [num_tokens, total_kv_heads * D] (sliced to the local kv-heads as at :256-258).
        Returns the pre-``o_proj`` activations [num_tokens, H * D]."""
        T = qkv.shape[0]
        D = self.head_dim
        q = qkv[:, :self.q_size].view(T, self.num_heads, D)
        k = qkv[:, self.q_size:self.q_size + self.kv_size].view(T, self.tp_kv_heads, D)
        v = qkv[:, self.q_size + self.kv_size:].view(T, self.tp_kv_heads, D)
        gk = gk.view(T, -1, D)
        if gk.shape[1] != self.tp_kv_heads:
            s = self.tp_kv_heads * self.tp_rank
            gk = gk[:, s:s + self.tp_kv_heads]
        # q, k, gk stay RAW: the feature map and the gate are applied inside the GLA kernels
        kv_cache = kv_caches.gla_cache
        idx = kv_caches.state_indices_tensor
        if getattr(attn_metadata, "num_prefills", 0) != 0:
            return self._prefill_and_mix_infer(q, k, v, gk, kv_cache, idx, attn_metadata)
        return self._decode(q, k, v, gk, kv_cache, idx, attn_metadata)
