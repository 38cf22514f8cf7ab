"""Attention backend of the meta-token SWA path -- the boundary vLLM calls.

Same class names, method names, argument meaning and error behaviour as the reference
vllm_hymeta/attention/meta_attention.py (Backend :53-126, Metadata :128-409, Builder :411-608,
Impl :611-982), re-implemented over the sm_100a kernels:

* KV write          -> ops.kv_write_paged            (was torch.ops._C_cache_ops.reshape_and_cache_flash)
* prefill           -> ops.swa_meta_prefill_varlen   (was 2x flash-attn + TorchScript merge)
* decode            -> ops.swa_meta_decode_paged     (was fwd_kvcache + flash-attn + merge)

The reference targets the vLLM 0.10.0 V0 engine ABCs (``vllm.attention.backends.abstract``), which
the vLLM installed next to this package no longer ships; the classes below therefore bind to those
ABCs only when they import, and are otherwise plain classes with the same surface.
"""
from __future__ import annotations

from dataclasses import dataclass
from itertools import accumulate
from typing import Any, Dict, List, Optional, Tuple, Type

import numpy as np
import torch

from ..slot_cache import PAD_SLOT_ID
from .fa_kernel_metatoken import (metatoken_flash_attn_varlen_func,
                                  metatoken_flash_attn_with_kvcache)

try:  # vLLM 0.10.0 (V0) ABCs, when present
    from vllm.attention.backends.abstract import (AttentionBackend, AttentionImpl,  # type: ignore
                                                  AttentionMetadata, AttentionType)
    _HAVE_V0 = True
except Exception:  # pragma: no cover - depends on the installed vLLM
    _HAVE_V0 = False
    AttentionBackend = AttentionImpl = object

    class AttentionType:  # same string constants as vLLM
        DECODER = "decoder"
        ENCODER = "encoder"
        ENCODER_ONLY = "encoder_only"
        ENCODER_DECODER = "encoder_decoder"

    @dataclass
    class AttentionMetadata:  # the three fields of vLLM's base dataclass the path reads
        num_prefills: int
        num_prefill_tokens: int
        num_decode_tokens: int
        slot_mapping: Optional[torch.Tensor]
        multi_modal_placeholder_index_maps: Optional[dict] = None
        enable_kv_scales_calculation: bool = True


class MetaAttentionBackend(AttentionBackend):

    accept_output_buffer: bool = True

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        # the reference lists flash-attn's sizes (:57-58); the sm_100a kernels are built for the
        # one head size every Hymeta model uses
        return [128]

    @staticmethod
    def get_name() -> str:
        return "META_FLASH_ATTN"

    @staticmethod
    def get_impl_cls() -> Type["MetaAttentionImpl"]:
        return MetaAttentionImpl

    @staticmethod
    def get_metadata_cls() -> Type["MetaAttentionMetadata"]:
        return MetaAttentionMetadata

    @staticmethod
    def get_builder_cls() -> Type["MetaAttentionMetadataBuilder"]:
        return MetaAttentionMetadataBuilder

    @staticmethod
    def get_state_cls():
        from vllm.attention.backends.utils import CommonAttentionState  # type: ignore
        return CommonAttentionState

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        if block_size % 16 != 0 or block_size < 16:
            raise ValueError(
                f"Invalid block size {block_size}. Must be a multiple of 16 and at least 16.")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def swap_blocks(src_kv_cache: torch.Tensor, dst_kv_cache: torch.Tensor,
                    src_to_dst: torch.Tensor) -> None:
        """Copy whole blocks between two caches (possibly host <-> device); ``src_to_dst`` is an
        [n, 2] int64 tensor of (src block, dst block) pairs (reference :94-111)."""
        pairs = src_to_dst.to("cpu").tolist()
        for s, d in pairs:
            dst_kv_cache[:, d].copy_(src_kv_cache[:, s], non_blocking=True)

    @staticmethod
    def copy_blocks(kv_caches: List[torch.Tensor], src_to_dists: torch.Tensor) -> None:
        """In-cache block copies for every layer (reference :113-126)."""
        if src_to_dists.numel() == 0:
            return
        src = src_to_dists[:, 0].to(torch.long)
        dst = src_to_dists[:, 1].to(torch.long)
        for kv in kv_caches:
            kv[:, dst.to(kv.device)] = kv[:, src.to(kv.device)]


@dataclass
class MetaAttentionMetadata(AttentionMetadata):
    """Field-for-field the reference dataclass (:128-190); encoder / cross-attention fields are
    omitted (decoder-only models)."""
    seq_lens: Optional[List[int]] = None
    seq_lens_tensor: Optional[torch.Tensor] = None
    max_prefill_seq_len: int = 0
    max_decode_seq_len: int = 0
    context_lens_tensor: Optional[torch.Tensor] = None
    block_tables: Optional[torch.Tensor] = None
    use_cuda_graph: bool = False
    max_query_len: Optional[int] = None
    max_decode_query_len: Optional[int] = None
    query_start_loc: Optional[torch.Tensor] = None
    seq_start_loc: Optional[torch.Tensor] = None
    _cached_prefill_metadata: Optional["MetaAttentionMetadata"] = None
    _cached_decode_metadata: Optional["MetaAttentionMetadata"] = None

    @property
    def prefill_metadata(self) -> Optional["MetaAttentionMetadata"]:
        """View of the first ``num_prefills`` requests / ``num_prefill_tokens`` tokens (:222-284)."""
        if self.num_prefills == 0:
            return None
        if self._cached_prefill_metadata is not None:
            return self._cached_prefill_metadata
        assert self.seq_lens is not None and self.seq_lens_tensor is not None
        n, t = self.num_prefills, self.num_prefill_tokens
        sl = lambda x, e: None if x is None else x[:e]
        self._cached_prefill_metadata = MetaAttentionMetadata(
            num_prefills=n, num_prefill_tokens=t, num_decode_tokens=0,
            slot_mapping=sl(self.slot_mapping, t),
            multi_modal_placeholder_index_maps=self.multi_modal_placeholder_index_maps,
            enable_kv_scales_calculation=self.enable_kv_scales_calculation,
            seq_lens=self.seq_lens[:n], seq_lens_tensor=self.seq_lens_tensor[:n],
            max_query_len=self.max_query_len, max_prefill_seq_len=self.max_prefill_seq_len,
            max_decode_query_len=0, max_decode_seq_len=0,
            query_start_loc=sl(self.query_start_loc, n + 1), seq_start_loc=sl(self.seq_start_loc, n + 1),
            context_lens_tensor=sl(self.context_lens_tensor, n), block_tables=sl(self.block_tables, n),
            use_cuda_graph=False)
        return self._cached_prefill_metadata

    @property
    def decode_metadata(self) -> Optional["MetaAttentionMetadata"]:
        """View of the decode requests, query_start_loc rebased to 0 (:286-336)."""
        if self.num_decode_tokens == 0:
            return None
        if self._cached_decode_metadata is not None:
            return self._cached_decode_metadata
        assert self.seq_lens_tensor is not None
        n, t = self.num_prefills, self.num_prefill_tokens
        sl = lambda x, s: None if x is None else x[s:]
        self._cached_decode_metadata = MetaAttentionMetadata(
            num_prefills=0, num_prefill_tokens=0, num_decode_tokens=self.num_decode_tokens,
            slot_mapping=sl(self.slot_mapping, t), multi_modal_placeholder_index_maps=None,
            enable_kv_scales_calculation=True, seq_lens=None,
            seq_lens_tensor=self.seq_lens_tensor[n:], max_decode_query_len=self.max_decode_query_len,
            max_query_len=self.max_query_len, max_prefill_seq_len=0,
            max_decode_seq_len=self.max_decode_seq_len,
            query_start_loc=(self.query_start_loc[n:] - self.query_start_loc[n])
            if self.query_start_loc is not None else None,
            seq_start_loc=sl(self.seq_start_loc, n), context_lens_tensor=None,
            block_tables=sl(self.block_tables, n), use_cuda_graph=self.use_cuda_graph)
        return self._cached_decode_metadata

    def advance_step(self, model_input, sampled_token_ids: Optional[torch.Tensor], block_size: int,
                     num_seqs: int, num_queries: int, turn_prefills_into_decodes: bool = False):
        """Update the metadata in place to advance one decode step (multi-step scheduling; reference
        :338-409, same assertions).  The device half -- seq_lens_tensor += 1, new slot_mapping through
        the block tables, -1 for CUDA-graph padding rows, next input tokens / positions -- is ONE
        launch of ``ops.advance_step`` (was ``ops.advance_step_flashattn``)."""
        from .. import ops
        if num_seqs != num_queries:
            assert num_seqs > num_queries
            assert self.use_cuda_graph
        if turn_prefills_into_decodes:
            assert self.num_decode_tokens + self.num_prefills == num_seqs
            self.num_decode_tokens += self.num_prefills
            self.num_prefills = 0
            self.num_prefill_tokens = 0
            self.max_prefill_seq_len = 0
            self.max_query_len = 1
            self.slot_mapping = self.slot_mapping[:num_seqs]
        else:
            assert self.seq_lens is not None
            assert self.max_decode_seq_len == max(self.seq_lens)
        assert self.num_prefills == 0
        assert self.num_prefill_tokens == 0
        assert self.num_decode_tokens == num_seqs
        assert self.slot_mapping.shape == (num_seqs, )
        assert self.seq_lens is not None
        assert len(self.seq_lens) == num_seqs
        assert self.seq_lens_tensor is not None
        assert self.seq_lens_tensor.shape == (num_seqs, )
        assert self.max_query_len == 1
        assert self.max_prefill_seq_len == 0
        assert self.query_start_loc is not None
        assert self.query_start_loc.shape == (num_queries + 1, )
        assert self.seq_start_loc is not None
        assert self.seq_start_loc.shape == (num_seqs + 1, )
        assert self.context_lens_tensor is not None
        assert self.context_lens_tensor.shape == (num_queries, )
        assert self.block_tables is not None
        assert self.block_tables.shape[0] == num_seqs
        # only the live queries advance: tensors may be padded to the captured batch size
        for i in range(num_queries):
            self.seq_lens[i] += 1
        self.max_decode_seq_len = max(self.seq_lens)
        self._cached_decode_metadata = None
        ops.advance_step(num_seqs, num_queries, block_size, self.seq_lens_tensor, self.slot_mapping,
                         self.block_tables,
                         input_tokens=getattr(model_input, "input_tokens", None),
                         sampled_token_ids=sampled_token_ids,
                         input_positions=getattr(model_input, "input_positions", None))


# ----------------------------------------------------------------------------------------------
# slot-mapping helpers of vLLM 0.10.0 (vllm/attention/backends/utils.py), restated: they are vLLM
# code the reference builder calls (:487-494) and the installed vLLM no longer ships them
# ----------------------------------------------------------------------------------------------
def compute_slot_mapping_start_idx(is_prompt: bool, query_len: int, context_len: int,
                                   sliding_window: Optional[int]) -> int:
    """Prompt tokens that fall out of the sliding window are never cached."""
    if is_prompt and sliding_window is not None:
        return max(0, query_len - sliding_window)
    return 0


def compute_slot_mapping(is_profile_run: bool, slot_mapping: List, seq_len: int,
                         context_len: int, start_idx: int, block_size: int,
                         block_table: Optional[List[int]]) -> None:
    """Appends ONE int64 numpy chunk per sequence to ``slot_mapping`` (a list of chunks that
    ``build`` concatenates): PAD for tokens that are never cached, else
    ``block_table[i // block_size] * block_size + i % block_size`` for i in [context, seq_len) --
    vectorised, a 8192-token prompt is one arange instead of 8192 Python iterations."""
    if is_profile_run:
        slot_mapping.append(np.full(seq_len - context_len, PAD_SLOT_ID, dtype=np.int64))
        return
    n_pad = max(0, start_idx - context_len)
    lo = max(start_idx, context_len)
    i = np.arange(lo, max(lo, seq_len), dtype=np.int64)
    bt = np.asarray(block_table, dtype=np.int64)
    chunk = bt[i // block_size] * block_size + i % block_size if i.size else i
    if n_pad:
        chunk = np.concatenate([np.full(n_pad, PAD_SLOT_ID, dtype=np.int64), chunk])
    slot_mapping.append(chunk)


@dataclass
class SeqData:
    """What the builder needs per sequence (the reference reads the same values from vLLM's
    ``InterDataForSeqGroup``, :435-494)."""
    is_prompt: bool
    seq_len: int                 # context + new tokens (for decode: possibly window-clipped)
    query_len: int
    context_len: int
    block_table: Optional[List[int]]
    curr_sliding_window_block: int = 0
    orig_seq_len: Optional[int] = None
    prefix_cache_hit: bool = False


class MetaAttentionMetadataBuilder:
    """Builds MetaAttentionMetadata with on-device tensors (reference :411-608).  ``add`` takes
    plain per-sequence data so the logic is independent of the vLLM version; ``_add_seq_group`` /
    ``build`` adapt the V0 ``input_builder`` protocol onto it when vLLM drives it."""

    def __init__(self, input_builder=None, *, block_size: Optional[int] = None,
                 sliding_window: Optional[int] = None, device="cuda",
                 chunked_prefill_enabled: bool = False, graph_block_tables=None,
                 pin_memory: bool = False):
        self.input_builder = input_builder
        if input_builder is not None:
            self.runner = input_builder.runner
            self.sliding_window = input_builder.sliding_window
            self.block_size = input_builder.block_size
            self.device = self.runner.device
            self.chunked_prefill_enabled = input_builder.chunked_prefill_enabled
            self.graph_block_tables = getattr(self.runner, "graph_block_tables", None)
            self.pin_memory = getattr(self.runner, "pin_memory", False)
        else:
            self.runner = None
            self.sliding_window = sliding_window
            self.block_size = block_size
            self.device = device
            self.chunked_prefill_enabled = chunked_prefill_enabled
            self.graph_block_tables = graph_block_tables
            self.pin_memory = pin_memory
        self.prepare()

    def prepare(self):
        self.slot_mapping: List[np.ndarray] = []     # one int64 chunk per sequence
        self.prefill_seq_lens: List[int] = []
        self.context_lens: List[int] = []
        self.block_tables: List[List[int]] = []
        self.curr_seq_lens: List[int] = []
        self.seq_lens: List[int] = []
        self.query_lens: List[int] = []
        self.num_prefills = 0
        self.num_prefill_tokens = 0
        self.num_decode_tokens = 0
        self.has_prefix_cache_hit = False

    def add(self, s: SeqData, prefix_cache_hit: bool = False) -> None:
        """Context length, block table (windowed for decode) and slot mapping of one sequence
        (:435-494)."""
        self.context_lens.append(s.context_len)
        self.seq_lens.append(s.seq_len)
        self.query_lens.append(s.query_len)
        if s.is_prompt:
            self.num_prefills += 1
            self.num_prefill_tokens += s.query_len
            self.prefill_seq_lens.append(s.orig_seq_len if s.orig_seq_len is not None else s.seq_len)
        else:
            self.num_decode_tokens += s.query_len
            self.curr_seq_lens.append(s.seq_len)
        block_table: List[int] = []
        if prefix_cache_hit:
            block_table = list(s.block_table)
        elif (self.chunked_prefill_enabled or not s.is_prompt) and s.block_table is not None:
            if s.curr_sliding_window_block == 0:
                block_table = list(s.block_table)
            else:
                block_table = list(s.block_table[-s.curr_sliding_window_block:])
        self.block_tables.append(block_table)
        is_profile_run = s.block_table is None or len(s.block_table) == 0
        seq_len = s.orig_seq_len if s.orig_seq_len is not None else s.seq_len
        start_idx = compute_slot_mapping_start_idx(s.is_prompt, s.query_len, s.context_len,
                                                   self.sliding_window)
        compute_slot_mapping(is_profile_run, self.slot_mapping, seq_len, s.context_len, start_idx,
                             self.block_size, s.block_table)

    def _add_seq_group(self, inter_data, chunked_prefill_enabled: bool, prefix_cache_hit: bool):
        self.chunked_prefill_enabled = chunked_prefill_enabled
        for (seq_id, seq_len, curr_seq_len, query_len, context_len, win_blocks) in zip(
                inter_data.seq_ids, inter_data.orig_seq_lens, inter_data.seq_lens,
                inter_data.query_lens, inter_data.context_lens,
                inter_data.curr_sliding_window_blocks):
            bt = None if inter_data.block_tables is None else inter_data.block_tables.get(seq_id)
            self.add(SeqData(inter_data.is_prompt, curr_seq_len, query_len, context_len, bt,
                             win_blocks, orig_seq_len=seq_len), prefix_cache_hit)

    def _get_graph_runner_block_tables(self, num_seqs: int, block_tables: List[List[int]]) -> torch.Tensor:
        max_batch_size, max_blocks = self.graph_block_tables.shape
        assert max_batch_size >= num_seqs
        g = self.graph_block_tables[:num_seqs]
        for i, bt in enumerate(block_tables):
            if bt:
                n = min(len(bt), max_blocks)      # look-ahead blocks beyond the graph width are unused
                g[i, :n] = bt[:n]
        return self._h2d(g if isinstance(g, np.ndarray) else g.numpy(), torch.int32)

    def _h2d(self, data, dtype):
        pin = self.pin_memory and torch.cuda.is_available()
        if isinstance(data, np.ndarray):
            t = torch.from_numpy(data).to(dtype)
            t = t.pin_memory() if pin else t
        else:
            t = torch.tensor(data, dtype=dtype, pin_memory=pin)
        return t.to(self.device, non_blocking=True)

    def build(self, seq_lens: Optional[List[int]] = None, query_lens: Optional[List[int]] = None,
              cuda_graph_pad_size: int = -1, batch_size: Optional[int] = None) -> MetaAttentionMetadata:
        if self.input_builder is not None and not self.seq_lens:
            hit = any(d.prefix_cache_hit for d in self.input_builder.inter_data_list)
            for d in self.input_builder.inter_data_list:
                self._add_seq_group(d, self.input_builder.chunked_prefill_enabled, hit)
        seq_lens = list(seq_lens if seq_lens is not None else self.seq_lens)
        query_lens = list(query_lens if query_lens is not None else self.query_lens)
        use_captured_graph = cuda_graph_pad_size != -1
        max_query_len = max(query_lens)
        decode_query_lens = query_lens[self.num_prefills:]
        max_decode_query_len = max(decode_query_lens) if decode_query_lens else 1
        max_prefill_seq_len = max(self.prefill_seq_lens, default=0)
        max_decode_seq_len = max(self.curr_seq_lens, default=0)
        num_decode_tokens = self.num_decode_tokens
        query_start_loc = list(accumulate(query_lens, initial=0))
        seq_start_loc = list(accumulate(seq_lens, initial=0))
        num_seqs = len(seq_lens)
        chunks = list(self.slot_mapping)
        if use_captured_graph:
            chunks.append(np.full(cuda_graph_pad_size, PAD_SLOT_ID, dtype=np.int64))
        slot_mapping = np.concatenate(chunks) if chunks else np.zeros(0, dtype=np.int64)
        if use_captured_graph:
            num_decode_tokens = batch_size - self.num_prefill_tokens
            block_tables = self._get_graph_runner_block_tables(num_seqs, self.block_tables)
        else:
            width = max((len(b) for b in self.block_tables), default=0)
            rows = [b + [0] * (width - len(b)) for b in self.block_tables]
            block_tables = self._h2d(np.asarray(rows, dtype=np.int32).reshape(len(rows), width), torch.int32)
        assert max_query_len > 0, "query_lens: {}".format(query_lens)
        return MetaAttentionMetadata(
            num_prefills=self.num_prefills, slot_mapping=self._h2d(slot_mapping, torch.long),
            num_prefill_tokens=self.num_prefill_tokens, num_decode_tokens=num_decode_tokens,
            seq_lens=seq_lens, multi_modal_placeholder_index_maps=None,
            enable_kv_scales_calculation=True, seq_lens_tensor=self._h2d(seq_lens, torch.int32),
            max_query_len=max_query_len, max_decode_query_len=max_decode_query_len,
            max_prefill_seq_len=max_prefill_seq_len, max_decode_seq_len=max_decode_seq_len,
            query_start_loc=self._h2d(query_start_loc, torch.int32),
            seq_start_loc=self._h2d(seq_start_loc, torch.int32),
            context_lens_tensor=self._h2d(self.context_lens, torch.int32),
            block_tables=block_tables, use_cuda_graph=use_captured_graph)


def get_num_prefill_decode_query_kv_tokens(attn_metadata) -> Tuple[int, int, int]:
    """(prefill query tokens, prefill kv tokens, decode query tokens) for decoder self-attention
    (vLLM 0.10.0 attention/backends/utils.py, called at reference :814-816)."""
    return (attn_metadata.num_prefill_tokens, attn_metadata.num_prefill_tokens,
            attn_metadata.num_decode_tokens)


class MetaAttentionImpl(AttentionImpl):
    """Token layout (reference :612-639):
    |<- num_meta_tokens (only while the meta caches warm up) ->|<- prefill tokens ->|<- decode tokens ->|
    """

    def __init__(self, num_heads: int, head_size: int, scale: float, num_kv_heads: int,
                 alibi_slopes: Optional[List[float]], sliding_window: Optional[int],
                 kv_cache_dtype: str, blocksparse_params: Optional[Dict[str, Any]] = None,
                 logits_soft_cap: Optional[float] = None, attn_type: str = AttentionType.DECODER,
                 kv_sharing_target_layer_name: Optional[str] = None, use_irope: bool = False,
                 num_meta_tokens: int = 128):
        if kv_sharing_target_layer_name is not None:
            raise NotImplementedError("KV sharing is not supported in V0 MetaAttention.")
        if blocksparse_params is not None:
            raise ValueError("MetaAttention does not support blocksparse attention.")
        if alibi_slopes is not None:
            raise NotImplementedError("MetaAttention (sm_100a) does not support alibi slopes.")
        if logits_soft_cap not in (None, 0, 0.0):
            raise NotImplementedError("MetaAttention (sm_100a) does not support logits soft cap.")
        if kv_cache_dtype not in ("auto", "bfloat16"):
            raise NotImplementedError(
                f"MetaAttention does not support {kv_cache_dtype} kv-cache on this device "
                "(bf16 KV only; no reference configuration uses fp8, SURVEY 2a).")
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("MetaAttention supports decoder self-attention only.")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_kv_heads
        self.alibi_slopes = None
        # (left, right) in flash-attn's convention; W-1 keys before the query (reference :671-672)
        self.sliding_window = ((sliding_window - 1, 0) if sliding_window is not None else (-1, -1))
        self.kv_cache_dtype = kv_cache_dtype
        self.num_meta_tokens = num_meta_tokens
        self.logits_soft_cap = 0
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        support_head_sizes = MetaAttentionBackend.get_supported_head_sizes()
        if self.head_size not in support_head_sizes:
            raise ValueError(f"Head size {head_size} is not supported by MetaAttention. "
                             f"Supported head sizes are {support_head_sizes}.")
        self.attn_type = attn_type

    def forward(self, layer, query: torch.Tensor, key: torch.Tensor, value: torch.Tensor,
                query2: Optional[torch.Tensor], key2: Optional[torch.Tensor],
                value2: Optional[torch.Tensor], kv_cache: torch.Tensor,
                attn_metadata: MetaAttentionMetadata, output: Optional[torch.Tensor] = None,
                output_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        """query [T,H,D]; key,value [T,Hkv,D]; query2 [M,H,D]|None; key2,value2 [M,Hkv,D]|None;
        kv_cache [2,nb,bs,Hkv,D] (numel 0 in the profiling run); output [(M)+T,H,D] written IN
        PLACE (reference :703-982)."""
        from .. import ops
        assert output is not None, "Output tensor must be provided."
        if output_scale is not None:
            raise NotImplementedError("fused output quantization is not supported in MetaAttention.")
        window_size = self.sliding_window
        (num_prefill_query_tokens, num_prefill_kv_tokens,
         num_decode_query_tokens) = get_num_prefill_decode_query_kv_tokens(attn_metadata)
        # store the new keys/values in the paged cache first (:790-799); skipped when the fused RoPE
        # prologue (ops.rope_kv_write_paged) already stored them.  The rows of single-token decode
        # requests are not written here: the decode kernel stores them itself (same slots, PAD
        # slots skipped) and attends to them from registers -- one launch less per layer-step.
        kv_pending = (kv_cache.numel() > 0 and key is not None and value is not None
                      and not getattr(layer, "kv_already_written", False))
        fuse_decode_write = (kv_pending and num_decode_query_tokens > 0
                             and (attn_metadata.max_decode_query_len or 1) == 1)
        slot_mapping = attn_metadata.slot_mapping.flatten() if kv_pending else None
        if kv_pending:
            n_w = num_prefill_kv_tokens if fuse_decode_write else key.shape[0]
            if n_w > 0:
                ops.kv_write_paged(key[:n_w], value[:n_w], kv_cache[0], kv_cache[1], slot_mapping[:n_w])
        num_meta_prefill_tokens = num_prefill_query_tokens
        if query2 is not None:  # meta tokens are not cached yet: they are forwarded together
            num_meta_prefill_tokens += self.num_meta_tokens
        decode_query = query[num_prefill_query_tokens:]
        query = query[:num_prefill_query_tokens]
        assert query.shape[0] == num_prefill_query_tokens, \
            f"query.shape: {query.shape}, num_prefill_query_tokens: {num_prefill_query_tokens}"
        assert decode_query.shape[0] == num_decode_query_tokens, \
            f"decode_query.shape: {decode_query.shape}, num_decode_query_tokens: {num_decode_query_tokens}"

        prefill_meta = attn_metadata.prefill_metadata
        if prefill_meta is not None or query2 is not None:
            # the kernel writes straight into `output` when its rows are [H, D]-contiguous (the
            # reference copies: output[:n] = res[0], :920 -- 7.5 GB of traffic per layer at config 2)
            pre_out = output[:num_meta_prefill_tokens]
            if pre_out.dim() != 3 or pre_out.stride(2) != 1:
                pre_out = None
            no_cache = (kv_cache.numel() == 0 or prefill_meta is None
                        or prefill_meta.block_tables is None or prefill_meta.block_tables.numel() == 0)
            if prefill_meta is None:                       # meta tokens only (warm-up pass)
                cu = torch.zeros(1, dtype=torch.int32, device=output.device)
                res = metatoken_flash_attn_varlen_func(
                    q1=query, q2=query2, k1=key[:0], k2=key2, v1=value[:0], v2=value2,
                    cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=0, max_seqlen_k=0,
                    num_meta_tokens=self.num_meta_tokens, softmax_scale=self.scale, causal=True,
                    window_size=window_size, out=pre_out)
            elif no_cache:
                # q and k are the prompt and have the same length (:840-884)
                res = metatoken_flash_attn_varlen_func(
                    q1=query, q2=query2, k1=key[:num_prefill_kv_tokens], k2=key2,
                    v1=value[:num_prefill_kv_tokens], v2=value2,
                    cu_seqlens_q=prefill_meta.seq_start_loc, cu_seqlens_k=prefill_meta.seq_start_loc,
                    max_seqlen_q=prefill_meta.max_prefill_seq_len,
                    max_seqlen_k=prefill_meta.max_prefill_seq_len,
                    num_meta_tokens=self.num_meta_tokens, softmax_scale=self.scale, causal=True,
                    window_size=window_size, out=pre_out)
            else:
                # prefix-enabled / chunked prefill: keys come from the paged cache (:886-919)
                assert prefill_meta.seq_lens is not None and prefill_meta.query_start_loc is not None
                res = metatoken_flash_attn_varlen_func(
                    q1=query, q2=query2, k1=kv_cache[0], k2=key2, v1=kv_cache[1], v2=value2,
                    cu_seqlens_q=prefill_meta.query_start_loc, cu_seqlens_k=prefill_meta.seq_start_loc,
                    max_seqlen_q=prefill_meta.max_query_len, max_seqlen_k=max(prefill_meta.seq_lens),
                    num_meta_tokens=self.num_meta_tokens, softmax_scale=self.scale, causal=True,
                    window_size=window_size, block_tables=prefill_meta.block_tables, out=pre_out)
            if pre_out is None:
                output[:num_meta_prefill_tokens] = res[0]
        decode_meta = attn_metadata.decode_metadata
        if decode_meta is not None:
            assert decode_meta.max_decode_query_len is not None
            if decode_meta.max_decode_query_len > 1:
                # varlen decode (speculative / multi-token): keys from the paged cache (:924-953)
                res = metatoken_flash_attn_varlen_func(
                    q1=decode_query, q2=None, k1=kv_cache[0], k2=key2, v1=kv_cache[1], v2=value2,
                    cu_seqlens_q=decode_meta.query_start_loc, cu_seqlens_k=decode_meta.seq_start_loc,
                    max_seqlen_q=decode_meta.max_decode_query_len,
                    max_seqlen_k=decode_meta.max_decode_seq_len,
                    num_meta_tokens=self.num_meta_tokens, softmax_scale=self.scale, causal=True,
                    window_size=window_size, block_tables=decode_meta.block_tables)
                output[num_meta_prefill_tokens:] = res[0]
            else:
                # one token per request: single-pass split-K paged kernel writes straight into
                # the output rows (:954-981)
                metatoken_flash_attn_with_kvcache(
                    q1=decode_query, q2=None, k2=key2, v2=value2, key_cache=kv_cache[0],
                    value_cache=kv_cache[1], cache_seqlens=decode_meta.seq_lens_tensor,
                    num_meta_tokens=self.num_meta_tokens, softmax_scale=self.scale, causal=True,
                    window_size=window_size, block_tables=decode_meta.block_tables,
                    out=output[num_meta_prefill_tokens:num_meta_prefill_tokens + num_decode_query_tokens],
                    new_key=key[num_prefill_kv_tokens:] if fuse_decode_write else None,
                    new_value=value[num_prefill_kv_tokens:] if fuse_decode_write else None,
                    new_slot_mapping=slot_mapping[num_prefill_kv_tokens:] if fuse_decode_write else None)
        return output
