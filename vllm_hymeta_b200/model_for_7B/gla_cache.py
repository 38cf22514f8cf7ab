"""Slot-state cache of the 7B (inter-layer hybrid) model -- same names and call contract as the
reference vllm_hymeta/model_for_7B/gla_cache.py:8-32.  ``_copy_cache`` copies SLOTS of every layer
(the reference iterates the layer dimension of a single tensor and indexes heads)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List

import torch

from ..slot_cache import PAD_SLOT_ID, NativeSlotAllocator


def _empty():
    return torch.empty(0)


@dataclass
class GLACacheParams:
    gla_cache: torch.Tensor = field(default_factory=_empty)
    state_indices_tensor: torch.Tensor = field(default_factory=_empty)

    def at_layer_idx(self, layer_idx):
        return GLACacheParams(self.gla_cache[layer_idx, ...], self.state_indices_tensor)


class GLACacheManager(NativeSlotAllocator):

    def __init__(self, dtype, cache_shape, device="cuda"):
        super().__init__(cache_shape[1])
        self._gla_cache = torch.zeros(size=cache_shape, dtype=dtype, device=device)

    @property
    def cache(self):
        return self._gla_cache

    def _copy_cache(self, from_index: int, to_index: int):
        from .. import ops
        ops.state_copy_slot(self._gla_cache, from_index, to_index)

    def clear_slots(self, slots: List[int]) -> None:
        """Zero slots explicitly (the prefill kernel already starts context-0 requests from a zero
        state on device -- ``GatedLinearAttention._prefill_and_mix_infer`` -- so the model's forward
        does not need this; kept for callers that want reference ``_clear_prefill_cache``
        semantics on the tensor itself, modeling_gla_swa.py:391-414)."""
        if not slots:
            return
        from .. import ops
        ops.state_zero_slots(self._gla_cache, self._to_device(slots))

    def _to_device(self, slots: List[int], stage_only: bool = False) -> torch.Tensor:
        """Slot indices -> int32 device tensor through PINNED staging memory: a copy from pageable
        memory makes the driver synchronise the stream first, which would serialise the host side of
        the next step behind the whole GPU step in flight."""
        dev = self._gla_cache.device
        pin = dev.type == "cuda"
        t = torch.tensor(slots, dtype=torch.int32, pin_memory=pin)
        return t if stage_only else t.to(dev, non_blocking=True)

    def current_run_tensors(self, **kwargs):
        if "seqlen_agnostic_capture_inputs" not in kwargs:
            slots = self.prepare(kwargs["request_ids_to_seq_ids"], kwargs["finished_requests_ids"],
                                 self._copy_cache)
            return self._gla_cache, self._to_device(slots)
        return kwargs["seqlen_agnostic_capture_inputs"]

    def copy_inputs_before_cuda_graphs(self, input_buffers, **kwargs):
        slots = self.prepare(kwargs["request_ids_to_seq_ids"], kwargs["finished_requests_ids"],
                             self._copy_cache)
        _, idx_buf = input_buffers["seqlen_agnostic_capture_inputs"]
        pad = idx_buf.numel() - len(slots)
        assert pad >= 0
        idx_buf.copy_(self._to_device(slots + [PAD_SLOT_ID] * pad, stage_only=True), non_blocking=True)

    def get_seqlen_agnostic_capture_inputs(self, batch_size):
        return self._gla_cache, torch.as_tensor([PAD_SLOT_ID] * batch_size, dtype=torch.int32,
                                                device=self._gla_cache.device)
