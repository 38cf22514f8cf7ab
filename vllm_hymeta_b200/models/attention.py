"""SWA-with-meta-prefix module glue -- the reference ``MetaAttention`` layer
(vllm_hymeta/models/attention.py:25-283) and the q/k/v handling of ``FlashAttention``
(vllm_hymeta/models/modeling_hymeta.py:747-815), on the sm_100a backend.

``MetaAttention`` owns the backend impl and the (vLLM-allocated) paged KV tensor and allocates the
``[(M)+T, H*D]`` output buffer the impl writes in place; ``FlashAttention`` splits the meta tokens
off the front of the batch while the shared meta caches warm up and stores their K/V in
``meta_fattn_cache`` (they never enter the paged cache, :776-792).  Projections and RoPE are
vLLM-side -- ``forward`` takes post-RoPE q/k/v.
"""
from __future__ import annotations

from contextlib import contextmanager
from typing import Optional

import torch

from ..attention.meta_attention import MetaAttentionBackend, MetaAttentionImpl
from .hymeta_cache import HymetaCacheParams


class HymetaRotaryEmbedding:
    """Reference ``HymetaRotaryEmbedding`` (modeling_hymeta.py:184-234): same constructor, same
    bf16 cos|sin table, ``forward(positions, query, key)`` rotating in place -- on the sm_100a
    kernel instead of ``ops.rotary_embedding``.  NeoX style, rotary_dim == head_size only (every
    reference config, :1181-1188)."""

    def __init__(self, head_size: int, rotary_dim: int, max_position: int, base: float,
                 is_neox_style: bool = True, cache_dtype: torch.dtype = torch.bfloat16,
                 fp32_math: bool = False):
        if not is_neox_style or rotary_dim != head_size:
            raise NotImplementedError("only NeoX-style full-head rotary is on the Hymeta path")
        if cache_dtype != torch.bfloat16:
            raise NotImplementedError("cos_sin_cache is bf16 on the Hymeta path (:1187)")
        self.head_size, self.rotary_dim = head_size, rotary_dim
        self.max_position_embeddings, self.base = max_position, base
        self.is_neox_style, self.cache_dtype = is_neox_style, cache_dtype
        # False: per-operation bf16 rounding (vLLM 0.10.0's op, what the reference runs on);
        # True: fp32 math with one rounding (the op of the vLLM 0.22 installed here, bit for bit)
        self.fp32_math = fp32_math
        from .. import ops
        self.cos_sin_cache = ops.rope_cos_sin_cache(head_size, max_position, base, dtype=cache_dtype)

    def forward(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor):
        """query [T, H*D], key [T, Hkv*D] (any row stride); rotated in place and returned."""
        from .. import ops
        if self.cos_sin_cache.device != positions.device:
            self.cos_sin_cache = self.cos_sin_cache.to(positions.device)
        D = self.head_size
        ops.rope_kv_write_paged(query.view(query.shape[0], -1, D), key.view(key.shape[0], -1, D),
                                None, self.cos_sin_cache, positions, fp32_math=self.fp32_math)
        return query, key

    __call__ = forward


# ---------------------------------------------------------------------------------------------
# forward context + the custom ops of the reference (models/attention.py:286-409)
# ---------------------------------------------------------------------------------------------
class ForwardContext:
    """The three fields of vLLM's ForwardContext the ops read (reference :361-366)."""

    def __init__(self, attn_metadata, no_compile_layers, virtual_engine: int = 0):
        self.attn_metadata = attn_metadata
        self.no_compile_layers = no_compile_layers
        self.virtual_engine = virtual_engine


_LAYERS: dict = {}          # layer_name -> MetaAttention (vLLM: compilation_config.static_forward_context)
_forward_context: Optional[ForwardContext] = None


@contextmanager
def set_forward_context(attn_metadata, virtual_engine: int = 0, layers: Optional[dict] = None):
    """Stand-alone equivalent of ``vllm.forward_context.set_forward_context`` for callers that run
    the layers without a vLLM engine (tests, bench.py)."""
    global _forward_context
    prev = _forward_context
    _forward_context = ForwardContext(attn_metadata, _LAYERS if layers is None else layers, virtual_engine)
    try:
        yield _forward_context
    finally:
        _forward_context = prev


def get_forward_context() -> ForwardContext:
    """This package's context when one is set, else vLLM's (the engine sets it around the model's
    forward; reference :361)."""
    if _forward_context is not None:
        return _forward_context
    from vllm.forward_context import get_forward_context as _vllm_ctx  # type: ignore
    return _vllm_ctx()


def _layer_kv_cache(layer, ctx: ForwardContext) -> torch.Tensor:
    kv = layer.kv_cache
    return kv[ctx.virtual_engine] if isinstance(kv, (list, tuple)) else kv


def _unified_meta_attention_with_output(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                                        output: torch.Tensor, layer_name: str,
                                        query2: Optional[torch.Tensor] = None,
                                        key2: Optional[torch.Tensor] = None,
                                        value2: Optional[torch.Tensor] = None) -> None:
    ctx = get_forward_context()
    layer = ctx.no_compile_layers[layer_name]
    layer.impl.forward(layer, query, key, value, query2, key2, value2, _layer_kv_cache(layer, ctx),
                       ctx.attn_metadata, output=output)


def _unified_meta_attention(query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                            layer_name: str, query2: Optional[torch.Tensor] = None,
                            key2: Optional[torch.Tensor] = None,
                            value2: Optional[torch.Tensor] = None) -> torch.Tensor:
    rows = query.shape[0] + (query2.shape[0] if query2 is not None else 0)
    output = torch.empty((rows,) + tuple(query.shape[1:]), dtype=query.dtype, device=query.device)
    _unified_meta_attention_with_output(query, key, value, output, layer_name, query2, key2, value2)
    return output


def _register_custom_ops() -> None:
    """``torch.ops.vllm.unified_meta_attention[_with_output]`` -- same op names and arguments as the
    reference (:317-409).  The fakes carry the query2/key2/value2 parameters the reference's
    ``unified_meta_attention_fake`` forgets (:343-349), and the functional op's fake
    accounts for the meta rows in its output shape.  Registering twice (re-import, or the reference
    plugin loaded next to this one) is tolerated."""
    try:
        torch.library.custom_op("vllm::unified_meta_attention_with_output",
                                _unified_meta_attention_with_output, mutates_args=["output"])

        @torch.library.register_fake("vllm::unified_meta_attention_with_output")
        def _(query, key, value, output, layer_name, query2=None, key2=None, value2=None):
            return None

        torch.library.custom_op("vllm::unified_meta_attention", _unified_meta_attention, mutates_args=[])

        @torch.library.register_fake("vllm::unified_meta_attention")
        def _(query, key, value, layer_name, query2=None, key2=None, value2=None):
            rows = query.shape[0] + (query2.shape[0] if query2 is not None else 0)
            return query.new_empty((rows,) + tuple(query.shape[1:]))
    except RuntimeError:      # already defined in this process
        pass


_register_custom_ops()


class MetaAttention:
    """Reference models/attention.py:25-283 without the vLLM config plumbing: same ``forward``
    arguments.  ``attn_metadata`` may be passed explicitly; when omitted it comes from the forward
    context, and with ``use_direct_call=False`` the call goes through
    ``torch.ops.vllm.unified_meta_attention_with_output`` exactly as in the reference (:242-245)."""

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 per_layer_sliding_window: Optional[int] = None, num_meta_tokens: int = 128,
                 kv_cache_dtype: str = "auto", prefix: str = "", use_direct_call: bool = True):
        self.num_heads = num_heads
        self.head_size = head_size
        self.num_kv_heads = num_kv_heads if num_kv_heads is not None else num_heads
        self.sliding_window = per_layer_sliding_window
        self.attn_backend = MetaAttentionBackend
        self.impl = MetaAttentionImpl(num_heads, head_size, scale, self.num_kv_heads, None,
                                      per_layer_sliding_window, kv_cache_dtype,
                                      num_meta_tokens=num_meta_tokens)
        self.use_output = self.attn_backend.accept_output_buffer
        self.layer_name = prefix
        self.use_direct_call = use_direct_call
        if prefix:
            if prefix in _LAYERS:
                raise ValueError(f"Duplicate layer name: {prefix}")      # reference :164-165
            _LAYERS[prefix] = self
        self.kv_cache = torch.tensor([])          # replaced by bind_kv_cache (reference :170-176)
        self.kv_already_written = False           # set per call by FlashAttention.forward_fused

    def bind_kv_cache(self, kv_cache: torch.Tensor) -> None:
        self.kv_cache = kv_cache

    def forward(self, query, key, value, query2=None, key2=None, value2=None, attn_metadata=None,
                output_shape: Optional[torch.Size] = None) -> torch.Tensor:
        rows = query.shape[0] + (query2.shape[0] if query2 is not None else 0)
        hidden = self.num_heads * self.head_size
        output = torch.empty(output_shape if output_shape is not None else (rows, hidden),
                             dtype=query.dtype, device=query.device)
        q = query.view(-1, self.num_heads, self.head_size)
        k = key.view(-1, self.num_kv_heads, self.head_size) if key is not None else None
        v = value.view(-1, self.num_kv_heads, self.head_size) if value is not None else None
        q2 = query2.view(-1, self.num_heads, self.head_size) if query2 is not None else None
        o3 = output.view(-1, self.num_heads, self.head_size)
        if self.use_direct_call:
            if attn_metadata is None:
                ctx = get_forward_context()
                attn_metadata, kv_cache = ctx.attn_metadata, _layer_kv_cache(self, ctx)
            else:
                kv_cache = self.kv_cache
            self.impl.forward(self, q, k, v, q2, key2, value2, kv_cache, attn_metadata, output=o3)
        else:
            torch.ops.vllm.unified_meta_attention_with_output(q, k, v, o3, self.layer_name, q2, key2, value2)
        return output.view(-1, hidden)


class FlashAttention:
    """q/k/v handling of the reference ``FlashAttention.forward`` (modeling_hymeta.py:747-815)."""

    def __init__(self, num_heads: int, num_kv_heads: int, head_dim: int, num_meta_tokens: int = 128,
                 sliding_window: Optional[int] = None, layer_idx: Optional[int] = None):
        self.num_heads, self.num_kv_heads, self.head_dim = num_heads, num_kv_heads, head_dim
        self.q_size, self.kv_size = num_heads * head_dim, num_kv_heads * head_dim
        self.num_meta_tokens = num_meta_tokens
        self.layer_idx = layer_idx
        self.scaling = head_dim ** -0.5
        self.attn = MetaAttention(num_heads, head_dim, self.scaling, num_kv_heads,
                                  per_layer_sliding_window=sliding_window,
                                  num_meta_tokens=num_meta_tokens)

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, kv_caches: HymetaCacheParams,
                attn_metadata, has_meta_cache: Optional[bool] = False,
                is_vllm_testing: bool = False) -> torch.Tensor:
        """q [T, H*D], k, v [T, Hkv*D] after RoPE; rows = [meta tokens (warm-up only) | prefill |
        decode].  Returns [T, H*D] with the meta rows first."""
        M = self.num_meta_tokens
        q1, k1, v1 = q, k, v
        q2 = k2 = v2 = None
        if (not has_meta_cache) and M > 0 and (not is_vllm_testing):
            q1, k1, v1 = q[M:], k[M:], v[M:]
            # only the pass that first sees the meta tokens needs their queries (:780)
            q2 = q[:M].reshape(M, self.num_heads, self.head_dim)
            kv_caches.meta_fattn_cache[0].copy_(k[:M].reshape(M, self.num_kv_heads, self.head_dim))
            kv_caches.meta_fattn_cache[1].copy_(v[:M].reshape(M, self.num_kv_heads, self.head_dim))
        if M > 0:
            k2 = kv_caches.meta_fattn_cache[0]
            v2 = kv_caches.meta_fattn_cache[1]
        return self.attn.forward(q1, k1, v1, query2=q2, key2=k2, value2=v2,
                                 attn_metadata=attn_metadata)

    def forward_fused(self, qkv: torch.Tensor, positions: torch.Tensor, kv_caches: HymetaCacheParams,
                      attn_metadata, rotary_emb: HymetaRotaryEmbedding,
                      has_meta_cache: Optional[bool] = False,
                      is_vllm_testing: bool = False) -> torch.Tensor:
        """Same result as ``rotary_emb(new_positions, q, k)`` + ``forward`` (modeling_hymeta.py:
        757-796, positions shifted as :1314-1324) with the prologue in ONE launch:
        RoPE of q and k in place inside ``qkv`` [T, (H+2Hkv)*D] (the raw ``qkv_proj`` output),
        rotated k + v stored into their pages, and on the warm-up pass the first M rows stored into
        ``meta_fattn_cache`` instead.  ``positions`` are the raw (unshifted) positions of the
        non-meta rows."""
        from .. import ops
        M, H, Hkv, D = self.num_meta_tokens, self.num_heads, self.num_kv_heads, self.head_dim
        T = qkv.shape[0]
        q, k, v = qkv.split([self.q_size, self.kv_size, self.kv_size], dim=-1)
        first = (not has_meta_cache) and M > 0 and (not is_vllm_testing)
        n_meta = M if first else 0
        kv_cache = self.attn.kv_cache
        paged = kv_cache.numel() > 0
        if rotary_emb.cos_sin_cache.device != qkv.device:
            rotary_emb.cos_sin_cache = rotary_emb.cos_sin_cache.to(qkv.device)
        ops.rope_kv_write_paged(
            q.view(T, H, D), k.view(T, Hkv, D), v.view(T, Hkv, D), rotary_emb.cos_sin_cache,
            positions, M, kv_cache[0] if paged else None, kv_cache[1] if paged else None,
            attn_metadata.slot_mapping.flatten() if paged else None,
            kv_caches.meta_fattn_cache[0] if first else None,
            kv_caches.meta_fattn_cache[1] if first else None, n_meta,
            fp32_math=rotary_emb.fp32_math)
        q2 = q[:M].reshape(M, H, D) if first else None
        k2 = kv_caches.meta_fattn_cache[0] if M > 0 else None
        v2 = kv_caches.meta_fattn_cache[1] if M > 0 else None
        self.attn.kv_already_written = paged
        try:
            return self.attn.forward(q[n_meta:], k[n_meta:], v[n_meta:], query2=q2, key2=k2,
                                     value2=v2, attn_metadata=attn_metadata)
        finally:
            self.attn.kv_already_written = False
