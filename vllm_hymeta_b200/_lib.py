"""ctypes binding of the C-ABI kernel library (include/hymeta_b200.h).

There is no fallback: if ``csrc/libhymeta_b200.so`` is missing the import of any
op raises, and every op raises if the library reports an error.  Build it with
``python -c "import __graft_entry__ as g; g.build()"`` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# HYMETA_B200_LIB: development override (A/B builds of the kernel library, tools/ab_swa_prefill.py)
LIB_PATH = os.environ.get("HYMETA_B200_LIB") or os.path.join(_HERE, "csrc", "libhymeta_b200.so")

_i, _i64, _f, _p, _sz = C.c_int, C.c_int64, C.c_float, C.c_void_p, C.c_size_t

# name -> (restype, argtypes); mirrors include/hymeta_b200.h one to one
SIGNATURES = {
    "hymeta_abi_version": (_i, []),
    "hymeta_status_string": (C.c_char_p, [_i]),
    "hymeta_last_cuda_error": (_i, []),
    "hymeta_launch_count": (_i64, []),
    "hymeta_gla_decode": (_i, [_p] * 7 + [_i] * 6 + [_i64] * 12 + [_i, _f, _p]),
    "hymeta_gla_prefill_varlen": (_i, [_p] * 10 + [_i] * 7 + [_i64] * 13 + [_i, _f, _i, _i, _p]),
    "hymeta_gla_prefill_workspace_bytes": (_sz, [_i] * 4),
    "hymeta_gla_prefill_balanced_workspace_bytes": (_sz, [_i] * 2),
    "hymeta_gla_prefill_varlen_seg": (_i, [_p] * 10 + [_i] * 7 + [_i64] * 13 + [_i, _f, _i, _i, _i, _p, _sz, _p]),
    "hymeta_gla_decode_gnorm": (_i, [_p] * 7 + [_i] * 6 + [_i64] * 12 + [_i, _f, _p, _f, _p]),
    "hymeta_gla_prefill_varlen_seg_gnorm": (_i, [_p] * 10 + [_i] * 7 + [_i64] * 13 +
                                            [_i, _f, _i, _i, _i, _p, _sz, _p, _f, _p]),
    "hymeta_gla_gate_70b": (_i, [_p] * 6 + [_i] * 5 + [_i64] * 10 + [_f, _p]),
    "hymeta_gla_gate_7b": (_i, [_p] * 6 + [_i] * 4 + [_i64] * 12 + [_f, _p]),
    "hymeta_gla_gate_7b_g": (_i, [_p, _p, _i, _i, _i] + [_i64] * 4 + [_f, _p]),
    "hymeta_gla_prefill_varlen_fused": (_i, [_p] * 10 + [_i] * 7 + [_i64] * 13 +
                                        [_i, _f, _i, _i, _i, _p, _sz, _p, _f, _i, _p, _p, _p]),
    "hymeta_gla_decode_gate7b": (_i, [_p] * 7 + [_i] * 6 + [_i64] * 12 + [_i, _f, _p, _f, _f, _p]),
    "hymeta_kv_write_paged": (_i, [_p] * 5 + [_i] * 4 + [_i64] * 10 + [_p]),
    "hymeta_rope_kv_write_paged": (_i, [_p] * 6 + [_i, _p, _i64] + [_p] * 5 + [_i] * 6 + [_i64] * 20 + [_i, _p]),
    "hymeta_swa_decode_workspace_bytes": (_sz, [_i] * 8),
    "hymeta_swa_meta_decode_paged": (_i, [_p] * 9 + [_sz] + [_i] * 8 + [_i64, _i, _f] + [_i64] * 14 + [_p]),
    "hymeta_swa_decode_ticket_bytes": (_sz, [_i] * 3),
    "hymeta_swa_meta_decode_paged_static": (_i, [_p] * 8 + [_i] * 8 + [_i64, _i, _f] + [_i64] * 14 + [_p]),
    "hymeta_swa_meta_decode_paged_fused": (_i, [_p] * 9 + [_sz] + [_i] * 8 + [_i64, _i, _f] + [_i64] * 14 +
                                           [_p, _p] + [_i64] * 4 + [_p, _p, _p]),
    "hymeta_swa_meta_prefill_varlen": (_i, [_p] * 10 + [_i64, _i] + [_i] * 9 + [_f] + [_i64] * 16 + [_p]),
    "hymeta_advance_step": (_i, [_i, _i, _i, _p, _p, _p, _p, _p, _p, _i64, _p]),
    "hymeta_state_zero_slots": (_i, [_p, _p, _i, _i, _i64, _i64, _i64, _p]),
    "hymeta_state_copy_slot": (_i, [_p, _i, _i, _i, _i64, _i64, _i64, _p]),
    "hymeta_selftest_umma": (_i, [_p]),
    "hymeta_branch_sumsq": (_i, [_p, _p, _p, _i, _i, _i64, _i64, _p]),
    "hymeta_branch_norm_merge": (_i, [_p] * 6 + [_i, _i, _i, _f, _i64, _i64, _i64, _p]),
    "hymeta_branch_merge_tp_buffer_bytes": (_sz, [_i, _i64]),
    "hymeta_branch_norm_merge_tp": (_i, [_p] * 5 + [_i] * 4 + [_p, _i64, C.c_uint32, _f, _i64, _i64, _i64, _p]),
    "hymeta_branch_norm_merge_tp_sumsq": (_i, [_p] * 5 + [_i] * 4 + [_p, _i64, C.c_uint32, _f, _i64, _i64, _i64,
                                              _p, _p, _i, _p]),
    "hymeta_branch_norm_merge_tp_graph": (_i, [_p] * 5 + [_i] * 4 + [_p, _i64, _p, _f, _i64, _i64, _i64,
                                              _p, _p, _i, _p]),
    "hymeta_swa_meta_prefill_varlen_sumsq": (_i, [_p] * 10 + [_i64, _i] + [_i] * 9 + [_f] + [_i64] * 16 + [_p, _p]),
    "hymeta_slot_mgr_create": (_p, [_i]),
    "hymeta_slot_mgr_destroy": (None, [_p]),
    "hymeta_slot_mgr_release": (_i, [_p, _p, _i]),
    "hymeta_slot_mgr_prepare": (_i, [_p, _p, _p, _i, _p, _i] + [_p] * 6),
    "hymeta_slot_mgr_free_count": (_i, [_p]),
    "hymeta_slot_mgr_free_list": (_i, [_p, _p, _i]),
    "hymeta_slot_mgr_slot_of": (_i, [_p, C.c_char_p, _i64]),
    "hymeta_slot_mgr_live_count": (_i, [_p]),
    "hymeta_slot_mgr_live": (_i, [_p, _p, _sz, _p, _p, _p, _i]),
}

_lib = None


class HymetaKernelError(RuntimeError):
    pass


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: the sm_100a kernel library is not built and there is no "
                "fallback path. Run `python -c 'import __graft_entry__ as g; g.build()'`.")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)      # AttributeError if the library lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        if h.hymeta_abi_version() != 6:
            raise ImportError("hymeta_b200 ABI version mismatch")
        _lib = h
    return _lib


def check(status: int, what: str) -> None:
    """Map a hymeta_status to the exception types of the reference boundary
    (SURVEY 8b): bad arguments -> ValueError, launch failures -> RuntimeError."""
    if status == 0:
        return
    msg = f"{what}: {lib().hymeta_status_string(status).decode()} (status {status})"
    if status in (-1, -2, -3):
        raise ValueError(msg)
    if status == -4:
        msg += f", cudaError {lib().hymeta_last_cuda_error()}"
    raise HymetaKernelError(msg)
