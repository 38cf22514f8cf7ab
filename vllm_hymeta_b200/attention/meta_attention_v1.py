"""vLLM V1 adapter of the meta-token SWA backend.

The reference backend (vllm_hymeta/attention/meta_attention.py) targets the vLLM 0.10.0 V0 engine
ABCs, which the vLLM installed next to this package (0.22) no longer ships.  This module exposes the
SAME kernels behind the V1 interfaces of the installed vLLM (``vllm.v1.attention.backend``):

* ``MetaAttentionV1Backend``          -- name / impl / builder / KV-cache shape ``(2, nb, bs, Hkv, D)``
  (the reference's layout, meta_attention.py:81-92)
* ``MetaAttentionV1MetadataBuilder``  -- ``CommonAttentionMetadata`` -> ``MetaAttentionV1Metadata``
  (what the reference builder derives from V0 sequence groups, :411-608: query / sequence start
  offsets, block tables, slot mapping, the decode / prefill split)
* ``MetaAttentionV1Impl``             -- ``do_kv_cache_update`` (reshape_and_cache_flash, :790-799)
  and ``forward`` (:814-981): single-token requests go to the split-K paged decode kernel, every
  other request to the paged-key prefill kernel (V1 has no separate "no cache" branch: keys are
  always read through the block table, chunked prefill and prefix hits included).

V1 passes no ``query2 / key2 / value2``: the shared meta K/V are read from the layer object
(``layer.meta_key`` / ``layer.meta_value`` [M, Hkv, D]; absent or None = no meta prefix, e.g. the 7B
model's SWA layers), and the one warm-up pass that needs meta QUERIES calls
``metatoken_flash_attn_varlen_func`` directly from the model code, as the reference does.

The classes bind to vLLM's ABCs when they import and are plain classes with the same surface
otherwise, so the logic is testable without a vLLM engine (tests/test_host_logic_cpu.py,
tests/test_gpu_modules.py).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, ClassVar, List, Optional, Tuple

import torch

try:  # the installed vLLM's V1 ABCs
    from vllm.v1.attention.backend import (AttentionBackend, AttentionCGSupport,  # type: ignore
                                           AttentionImpl, AttentionMetadataBuilder, AttentionType,
                                           MultipleOf)
    _HAVE_V1 = True
except Exception:  # pragma: no cover - depends on the installed vLLM
    _HAVE_V1 = False
    AttentionBackend = AttentionImpl = object

    class AttentionMetadataBuilder:  # minimal stand-in with the V1 constructor
        def __init__(self, kv_cache_spec, layer_names, vllm_config, device):
            self.kv_cache_spec, self.layer_names = kv_cache_spec, layer_names
            self.vllm_config, self.device = vllm_config, device

        def __class_getitem__(cls, item):
            return cls

    class AttentionType:
        DECODER = "decoder"

    class AttentionCGSupport:
        UNIFORM_SINGLE_TOKEN_DECODE = 2

    class MultipleOf:
        def __init__(self, base: int):
            self.base = base


@dataclass
class MetaAttentionV1Metadata:
    """Batch layout (V1 reorders the batch so single-token requests come first,
    ``reorder_batch_threshold = 1``):  | decode tokens (1 per request) | prefill / extend tokens |"""
    num_actual_tokens: int
    num_decodes: int
    num_prefills: int
    num_prefill_tokens: int
    max_query_len: int
    max_seq_len: int
    query_start_loc: torch.Tensor            # [R+1] int32
    seq_lens: torch.Tensor                   # [R]   int32, context + new tokens
    block_table: torch.Tensor                # [R, max_blocks] int32
    slot_mapping: torch.Tensor               # [T]   int64
    prefill_query_start_loc: Optional[torch.Tensor] = None   # [P+1] int32, rebased to 0
    prefill_seq_start_loc: Optional[torch.Tensor] = None     # [P+1] int32, cumulative seq_lens
    max_prefill_query_len: int = 0
    causal: bool = True


class MetaAttentionV1Backend(AttentionBackend):
    accept_output_buffer: bool = True
    supported_dtypes: ClassVar[List[torch.dtype]] = [torch.bfloat16]
    supported_kv_cache_dtypes: ClassVar[List[str]] = ["auto", "bfloat16"]
    forward_includes_kv_cache_update: bool = False

    @staticmethod
    def get_name() -> str:
        return "CUSTOM" if _HAVE_V1 else "META_FLASH_ATTN"   # AttentionBackendEnum.CUSTOM slot

    @staticmethod
    def get_supported_kernel_block_sizes():
        return [MultipleOf(16)]

    @classmethod
    def get_supported_head_sizes(cls) -> List[int]:
        return [128]

    @staticmethod
    def get_impl_cls():
        return MetaAttentionV1Impl

    @staticmethod
    def get_builder_cls():
        return MetaAttentionV1MetadataBuilder

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int, head_size: int,
                           cache_dtype_str: str = "auto") -> Tuple[int, ...]:
        if block_size % 16 != 0 or block_size < 16:
            raise ValueError(
                f"Invalid block size {block_size}. Must be a multiple of 16 and at least 16.")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @classmethod
    def supports_compute_capability(cls, capability) -> bool:
        major = getattr(capability, "major", None)
        return major is None or major == 10          # sm_100a kernels only

    @classmethod
    def supports_attn_type(cls, attn_type: str) -> bool:
        return attn_type == AttentionType.DECODER


class MetaAttentionV1MetadataBuilder(AttentionMetadataBuilder):
    _cudagraph_support = AttentionCGSupport.UNIFORM_SINGLE_TOKEN_DECODE
    reorder_batch_threshold: Optional[int] = 1

    def __init__(self, kv_cache_spec, layer_names, vllm_config, device):
        super().__init__(kv_cache_spec, layer_names, vllm_config, device)
        self.block_size = getattr(kv_cache_spec, "block_size", None)

    def build(self, common_prefix_len: int, common_attn_metadata, fast_build: bool = False
              ) -> MetaAttentionV1Metadata:
        m = common_attn_metadata
        R, T = int(m.num_reqs), int(m.num_actual_tokens)
        qsl = m.query_start_loc[:R + 1].to(torch.int32)
        seq_lens = m.seq_lens[:R].to(torch.int32)
        qsl_cpu = m.query_start_loc_cpu[:R + 1]
        qlens = (qsl_cpu[1:] - qsl_cpu[:-1]).tolist()
        nd = 0                                           # leading single-token requests
        while nd < R and qlens[nd] == 1:
            nd += 1
        md = MetaAttentionV1Metadata(
            num_actual_tokens=T, num_decodes=nd, num_prefills=R - nd,
            num_prefill_tokens=T - nd, max_query_len=int(m.max_query_len),
            max_seq_len=int(m.max_seq_len), query_start_loc=qsl, seq_lens=seq_lens,
            block_table=m.block_table_tensor[:R], slot_mapping=m.slot_mapping[:T],
            causal=bool(getattr(m, "causal", True)))
        if R > nd:
            md.prefill_query_start_loc = qsl[nd:] - nd   # decode requests own one token each
            z = torch.zeros(1, dtype=torch.int32, device=seq_lens.device)
            md.prefill_seq_start_loc = torch.cat([z, torch.cumsum(seq_lens[nd:], 0, dtype=torch.int32)])
            md.max_prefill_query_len = max(qlens[nd:])
        return md

    def build_for_cudagraph_capture(self, common_attn_metadata) -> MetaAttentionV1Metadata:
        return self.build(0, common_attn_metadata)


class MetaAttentionV1Impl(AttentionImpl):

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: Optional[int] = None,
                 alibi_slopes: Optional[List[float]] = None, sliding_window: Optional[int] = None,
                 kv_cache_dtype: str = "auto", logits_soft_cap: Optional[float] = None,
                 attn_type: str = AttentionType.DECODER,
                 kv_sharing_target_layer_name: Optional[str] = None, **extra: Any) -> None:
        # same refusals as the V0 impl (reference meta_attention.py:640-701)
        if kv_sharing_target_layer_name is not None:
            raise NotImplementedError("KV sharing is not supported in MetaAttention.")
        if alibi_slopes is not None:
            raise NotImplementedError("MetaAttention (sm_100a) does not support alibi slopes.")
        if logits_soft_cap not in (None, 0, 0.0):
            raise NotImplementedError("MetaAttention (sm_100a) does not support logits soft cap.")
        if kv_cache_dtype not in ("auto", "bfloat16"):
            raise NotImplementedError(f"MetaAttention does not support {kv_cache_dtype} kv-cache (bf16 only).")
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("MetaAttention supports decoder self-attention only.")
        if head_size != 128:
            raise ValueError(f"Head size {head_size} is not supported by MetaAttention. "
                             "Supported head sizes are [128].")
        self.num_heads, self.head_size, self.scale = num_heads, head_size, float(scale)
        self.num_kv_heads = num_kv_heads if num_kv_heads is not None else num_heads
        # keys strictly before the query that stay visible (reference :671-672)
        self.window_left = sliding_window - 1 if sliding_window is not None else -1
        self.sliding_window = (self.window_left, 0) if sliding_window is not None else (-1, -1)
        self.kv_cache_dtype = kv_cache_dtype
        self.attn_type = attn_type
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads

    def do_kv_cache_update(self, layer, key: torch.Tensor, value: torch.Tensor, kv_cache: torch.Tensor,
                           slot_mapping: torch.Tensor) -> None:
        """``reshape_and_cache_flash`` (reference :790-799).  key/value may be padded beyond
        ``slot_mapping`` (CUDA-graph padding): the slot mapping's length is the token count."""
        from .. import ops
        n = slot_mapping.numel()
        k = key.view(-1, self.num_kv_heads, self.head_size)[:n]
        v = value.view(-1, self.num_kv_heads, self.head_size)[:n]
        ops.kv_write_paged(k, v, kv_cache[0], kv_cache[1], slot_mapping)

    def forward(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                kv_cache: torch.Tensor, attn_metadata: Optional[MetaAttentionV1Metadata],
                output: Optional[torch.Tensor] = None, output_scale: Optional[torch.Tensor] = None,
                output_block_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        """query [T(padded), H, D]; kv_cache [2, nb, bs, Hkv, D] already holding this step's K/V
        (``do_kv_cache_update`` ran first); output [T(padded), H, D] written in place."""
        from .. import ops
        assert output is not None, "Output tensor must be provided."
        if output_scale is not None or output_block_scale is not None:
            raise NotImplementedError("fused output quantization is not supported in MetaAttention.")
        if attn_metadata is None:                        # profiling run
            return output.fill_(0)
        md = attn_metadata
        H, D = self.num_heads, self.head_size
        q = query.view(-1, H, D)
        out = output.view(-1, H, D)
        mk = getattr(layer, "meta_key", None)
        mv = getattr(layer, "meta_value", None)
        kc, vc = kv_cache[0], kv_cache[1]
        nd, T = md.num_decodes, md.num_actual_tokens
        if nd > 0:
            ops.swa_meta_decode_paged(q[:nd], kc, vc, md.block_table[:nd], md.seq_lens[:nd], mk, mv,
                                      self.window_left, scale=self.scale, out=out[:nd])
        if md.num_prefills > 0:
            ops.swa_meta_prefill_varlen(q[nd:T], kc, vc, None, mk, mv, md.prefill_query_start_loc,
                                        md.prefill_seq_start_loc, md.max_prefill_query_len,
                                        self.window_left, scale=self.scale,
                                        block_tables=md.block_table[nd:], out=out[nd:T])
        return output


def register_v1_backend() -> bool:
    """Put the backend in vLLM's ``AttentionBackendEnum.CUSTOM`` slot (vLLM >= 0.11 registry).
    Returns False when the installed vLLM has no such registry."""
    try:
        from vllm.v1.attention.backends.registry import (AttentionBackendEnum,  # type: ignore
                                                         register_backend)
    except Exception:
        return False
    register_backend(AttentionBackendEnum.CUSTOM,
                     "vllm_hymeta_b200.attention.meta_attention_v1.MetaAttentionV1Backend")
    return True
