"""Constant-size slot cache: request/sequence -> state-slot bookkeeping.

Re-implements the behaviour of vLLM 0.10.0's ``ConstantSizeCache`` that the reference's cache
managers inherit from (reference vllm_hymeta/models/hymeta_cache.py:24-63,
vllm_hymeta/model_for_7B/gla_cache.py:17-32) -- that base class is vLLM code, not in the
reference tree, and no longer exists in the installed vLLM.  Pure host logic; device work (slot
zero-fill, slot copy) goes through the C-ABI kernels.
"""
from __future__ import annotations

from array import array
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


class NativeSlotAllocator:
    """Same contract as :class:`SlotAllocator`, bookkeeping in the C++ manager behind the C ABI
    (``hymeta_slot_mgr_*``, csrc/slot_mgr.cu).  ``prepare_batch`` resolves a whole run
    in one native call and returns the fork copies and the freshly admitted slots so the caller
    batches the device work (one zero-fill launch per batch instead of one per request per layer,
    reference modeling_hymeta.py:1202-1227)."""

    def __init__(self, max_batch_size: int):
        import ctypes as C
        from ._lib import lib
        self._C, self._lib = C, lib()
        self.max_batch_size = max_batch_size
        self.last_new_slots: List[int] = []      # slots admitted by the latest prepare()
        self._h = self._lib.hymeta_slot_mgr_create(max_batch_size)
        if not self._h:
            raise ValueError("max_batch_size must be >= 0")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.hymeta_slot_mgr_destroy(h)
            except Exception:          # interpreter shutdown: ctypes may already be torn down
                pass

    @staticmethod
    def _ids(ids) -> bytes:
        """Packed id list: NUL-terminated strings back to back (one encode for the batch)."""
        return ("\0".join(ids) + "\0").encode() if ids else b"\0"

    def prepare_batch(self, request_ids_to_seq_ids: Dict[str, List[int]],
                      finished_requests_ids: List[str]):
        """-> (slots, [(from, to) forks in order], [new slots])."""
        C = self._C
        reqs = [r for r, seqs in request_ids_to_seq_ids.items() for _ in seqs]
        seqs = array("q", [s for ss in request_ids_to_seq_ids.values() for s in ss])
        n = len(reqs)
        out = array("i", bytes(16 * max(n, 1)))           # slots | copy_from | copy_to | new
        base, _ = out.buffer_info()
        slots, cf, ct, new = (base + 4 * i * max(n, 1) for i in range(4))
        nc, nn = C.c_int(0), C.c_int(0)
        st = self._lib.hymeta_slot_mgr_prepare(
            self._h, self._ids(reqs), seqs.buffer_info()[0], n,
            self._ids(finished_requests_ids), len(finished_requests_ids), slots, cf, ct,
            C.byref(nc), new, C.byref(nn))
        m = max(n, 1)
        slots, cf, ct, new = (out[i * m:(i + 1) * m] for i in range(4))
        if st == -6:
            # pairs resolved before the free list ran dry stay assigned (as in the reference):
            # hand their forks / new slots to the caller with the error
            e = RuntimeError("no free state slot: more live sequences than max_num_seqs")
            e.copies, e.new_slots = list(zip(cf[:nc.value], ct[:nc.value])), list(new[:nn.value])
            raise e
        if st != 0:
            raise ValueError(f"hymeta_slot_mgr_prepare: status {st}")
        return (slots[:n].tolist(), list(zip(cf[:nc.value], ct[:nc.value])), new[:nn.value].tolist())

    def prepare(self, request_ids_to_seq_ids: Dict[str, List[int]], finished_requests_ids: List[str],
                copy_fn) -> List[int]:
        try:
            slots, copies, self.last_new_slots = self.prepare_batch(request_ids_to_seq_ids,
                                                                    finished_requests_ids)
        except RuntimeError as e:
            for src, dst in getattr(e, "copies", []):
                copy_fn(src, dst)
            self.last_new_slots = getattr(e, "new_slots", [])
            raise
        for src, dst in copies:
            copy_fn(src, dst)
        return slots

    def release(self, finished_requests_ids: List[str]) -> None:
        self._lib.hymeta_slot_mgr_release(self._h, self._ids(finished_requests_ids),
                                          len(finished_requests_ids))

    def slot_of(self, req_id: str, seq_id: int) -> int:
        s = self._lib.hymeta_slot_mgr_slot_of(self._h, req_id.encode(), seq_id)
        if s < 0:
            raise KeyError((req_id, seq_id))
        return s

    @property
    def free_cache_indices(self) -> List[int]:
        n = self._lib.hymeta_slot_mgr_free_count(self._h)
        buf = (self._C.c_int32 * max(n, 1))()
        self._lib.hymeta_slot_mgr_free_list(self._h, buf, n)
        return list(buf[:n])

    def live(self) -> List[Tuple[str, int, int]]:
        C = self._C
        need = C.c_size_t(0)
        n = self._lib.hymeta_slot_mgr_live(self._h, None, 0, C.byref(need), None, None, 0)
        ids = C.create_string_buffer(max(need.value, 1))
        seqs, slots = (C.c_int64 * max(n, 1))(), (C.c_int32 * max(n, 1))()
        self._lib.hymeta_slot_mgr_live(self._h, ids, need.value, C.byref(need), seqs, slots, n)
        names = ids.raw[:need.value].split(b"\0")[:n]
        return [(names[i].decode(), int(seqs[i]), int(slots[i])) for i in range(n)]

    @property
    def cache_indices_mapping(self) -> Dict[str, Dict[int, int]]:
        """Snapshot in the shape the model reads (modeling_hymeta.py:1206-1208)."""
        out: Dict[str, Dict[int, int]] = {}
        for r, s, sl in self.live():
            out.setdefault(r, {})[s] = sl
        return out
