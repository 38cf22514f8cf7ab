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
        ops.state_zero_slots(self._gla_cache, torch.tensor(slots, dtype=torch.int32,
                                                           device=self._gla_cache.device))

    def current_run_tensors(self, **kwargs):
        if "seqlen_agnostic_capture_inputs" not in kwargs:
            slots = self.prepare(kwargs["request_ids_to_seq_ids"], kwargs["finished_requests_ids"],
                                 self._copy_cache)
            idx = torch.as_tensor(slots, dtype=torch.int32, device=self._gla_cache.device)
            return self._gla_cache, idx
        return kwargs["seqlen_agnostic_capture_inputs"]

    def copy_inputs_before_cuda_graphs(self, input_buffers, **kwargs):
        slots = self.prepare(kwargs["request_ids_to_seq_ids"], kwargs["finished_requests_ids"],
                             self._copy_cache)
        _, idx_buf = input_buffers["seqlen_agnostic_capture_inputs"]
        pad = idx_buf.numel() - len(slots)
        assert pad >= 0
        idx_buf.copy_(torch.as_tensor(slots + [PAD_SLOT_ID] * pad, dtype=torch.int32),
                      non_blocking=True)

    def get_seqlen_agnostic_capture_inputs(self, batch_size):
        return self._gla_cache, torch.as_tensor([PAD_SLOT_ID] * batch_size, dtype=torch.int32,
                                                device=self._gla_cache.device)
