"""Slot-state caches of the 70B (intra-layer hybrid) model -- same names and call contract as the
reference vllm_hymeta/models/hymeta_cache.py:9-63, B200-native storage:

* ``hymeta_cache``      [L, max_num_seqs, Hs, D, D] bf16, ZERO-initialised (the reference uses
  ``torch.empty`` and then host-syncs on ``isnan()`` per request per layer).  ``Hs`` is
  the number of states per request: the reference geometry passes ``num_heads/tp`` (one state per
  q-head); passing ``num_kv_heads/tp`` stores one state per GQA group -- bit-identical contents,
  7x less HBM traffic.  The kernels accept either.
* ``meta_linear_cache`` [L, Hs, D, D]          shared GLA state after the meta tokens
* ``meta_fattn_cache``  [L, 2, M, Hkv, D]      shared meta K/V
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import torch

from ..slot_cache import PAD_SLOT_ID, NativeSlotAllocator


def _empty():
    return torch.empty(0)


@dataclass
class HymetaCacheParams:
    hymeta_cache: torch.Tensor = field(default_factory=_empty)
    meta_linear_cache: torch.Tensor = field(default_factory=_empty)
    meta_fattn_cache: torch.Tensor = field(default_factory=_empty)
    # slots of the current run, batch order [prefill requests | decode requests]
    state_indices_tensor: torch.Tensor = field(default_factory=_empty)

    def at_layer_idx(self, layer_idx):
        return HymetaCacheParams(self.hymeta_cache[layer_idx, ...],
                                 self.meta_linear_cache[layer_idx, ...],
                                 self.meta_fattn_cache[layer_idx, ...],
                                 self.state_indices_tensor)


class HymetaCacheManager(NativeSlotAllocator):

    def __init__(self, dtype, cache_shape, meta_linear_cache_shape, meta_fattn_cache_shape,
                 device="cuda"):
        super().__init__(cache_shape[1])          # max_batch_size is cache_shape[1]
        hymeta_cache = torch.zeros(size=cache_shape, dtype=dtype, device=device)
        meta_linear_cache = torch.zeros(size=meta_linear_cache_shape, dtype=dtype, device=device)
        meta_fattn_cache = torch.zeros(size=meta_fattn_cache_shape, dtype=dtype, device=device)
        self._hymeta_cache = (hymeta_cache, meta_linear_cache, meta_fattn_cache)
        self._device = device

    @property
    def cache(self):
        return self._hymeta_cache

    def _copy_cache(self, from_index: int, to_index: int):
        """Fork a request's state into another slot (all layers, one launch)."""
        from .. import ops
        ops.state_copy_slot(self._hymeta_cache[0], from_index, to_index)

    def _to_device(self, slots: List[int], stage_only: bool = False) -> torch.Tensor:
        """Slot indices -> int32 device tensor through PINNED staging memory (a copy from pageable
        memory makes the driver synchronise the stream first, serialising the host side of the next
        step behind the GPU step in flight)."""
        dev = self._hymeta_cache[0].device
        t = torch.tensor(slots, dtype=torch.int32, pin_memory=dev.type == "cuda")
        return t if stage_only else t.to(dev, non_blocking=True)

    def clear_slots(self, slots: List[int]) -> None:
        """Zero the states of newly admitted requests (reference ``_clear_prefill_cache``,
        modeling_hymeta.py:1202-1227) with one launch over all layers."""
        if not slots:
            return
        from .. import ops
        state = self._hymeta_cache[0]
        ops.state_zero_slots(state, self._to_device(slots))

    def current_run_tensors(self, **kwargs):
        """-> (state, meta_linear, meta_fattn, state_indices[B] int32).  Outside CUDA graphs the
        slots are resolved from ``request_ids_to_seq_ids`` / ``finished_requests_ids``; inside a
        captured graph the runner passes the capture buffers (``seqlen_agnostic_capture_inputs``)
        whose indices ``copy_inputs_before_cuda_graphs`` refreshed."""
        if "seqlen_agnostic_capture_inputs" not in kwargs:
            slots = self.prepare(kwargs["request_ids_to_seq_ids"], kwargs["finished_requests_ids"],
                                 self._copy_cache)
            idx = self._to_device(slots)
            c = self._hymeta_cache
        else:
            c, idx = kwargs["seqlen_agnostic_capture_inputs"]
        return (c[0], c[1], c[2], idx)

    def copy_inputs_before_cuda_graphs(self, input_buffers, **kwargs):
        """Refresh the captured slot-index buffer before a graph replay; rows beyond the live batch
        get PAD_SLOT_ID so the kernels skip them."""
        slots = self.prepare(kwargs["request_ids_to_seq_ids"], kwargs["finished_requests_ids"],
                             self._copy_cache)
        _, idx_buf = input_buffers["seqlen_agnostic_capture_inputs"]
        pad = idx_buf.numel() - len(slots)
        assert pad >= 0, "captured batch smaller than the live batch"
        idx_buf.copy_(self._to_device(slots + [PAD_SLOT_ID] * pad, stage_only=True),
                      non_blocking=True)

    def get_seqlen_agnostic_capture_inputs(self, batch_size):
        return self._hymeta_cache, torch.as_tensor([PAD_SLOT_ID] * batch_size, dtype=torch.int32,
                                                   device=self._hymeta_cache[0].device)
