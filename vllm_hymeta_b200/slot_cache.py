"""Constant-size slot cache: request/sequence -> state-slot bookkeeping.

Re-implements the behaviour of vLLM 0.10.0's ``ConstantSizeCache`` that the reference's cache
managers inherit from (reference vllm_hymeta/models/hymeta_cache.py:24-63,
vllm_hymeta/model_for_7B/gla_cache.py:17-32) -- that base class is vLLM code, not in the
reference tree, and no longer exists in the installed vLLM.  Pure host logic; device work (slot
zero-fill, slot copy) goes through the C-ABI kernels.
"""
from __future__ import annotations

from typing import Dict, List, Tuple

PAD_SLOT_ID = -1      # vllm.attention.backends.utils.PAD_SLOT_ID


class SlotAllocator:
    """Free-list over ``max_batch_size`` slots with the (request_id -> {seq_id -> slot}) map the
    model reads as ``cache_indices_mapping`` (reference modeling_hymeta.py:1206-1208)."""

    def __init__(self, max_batch_size: int):
        self.max_batch_size = max_batch_size
        self.cache_indices_mapping: Dict[str, Dict[int, int]] = {}
        self.free_cache_indices: List[int] = list(range(max_batch_size))

    # -- one (request, sequence) ------------------------------------------------------------
    def _assign(self, req_id: str, seq_id: int, finished: List[str], copy_fn) -> int:
        if req_id in finished:
            return PAD_SLOT_ID                       # finished before it ever ran
        if req_id not in self.cache_indices_mapping:
            if not self.free_cache_indices:
                raise RuntimeError("no free state slot: more live sequences than max_num_seqs")
            slot = self.free_cache_indices.pop()
            self.cache_indices_mapping[req_id] = {seq_id: slot}
            return slot
        seqs = self.cache_indices_mapping[req_id]
        if seq_id not in seqs:
            # a sibling sequence of a running request (parallel sampling n>1 / beam): fork the
            # state of an existing sibling into a fresh slot
            if not self.free_cache_indices:
                raise RuntimeError("no free state slot: more live sequences than max_num_seqs")
            slot = self.free_cache_indices.pop()
            src = next(iter(seqs.values()))
            copy_fn(src, slot)
            seqs[seq_id] = slot
            return slot
        return seqs[seq_id]

    def prepare(self, request_ids_to_seq_ids: Dict[str, List[int]], finished_requests_ids: List[str],
                copy_fn) -> List[int]:
        """Release finished requests, then return the slot of every (request, seq) of this run in
        batch order."""
        self.release(finished_requests_ids)
        return [self._assign(r, s, finished_requests_ids, copy_fn)
                for r, seqs in request_ids_to_seq_ids.items() for s in seqs]

    def release(self, finished_requests_ids: List[str]) -> None:
        for r in finished_requests_ids:
            if r in self.cache_indices_mapping:
                for slot in self.cache_indices_mapping.pop(r).values():
                    self.free_cache_indices.append(slot)

    def slot_of(self, req_id: str, seq_id: int) -> int:
        return self.cache_indices_mapping[req_id][seq_id]

    def live(self) -> List[Tuple[str, int, int]]:
        return [(r, s, sl) for r, m in self.cache_indices_mapping.items() for s, sl in m.items()]
