"""CPU-side checks of the C-ABI boundary: the library loads without a GPU/driver and exports
every symbol include/hymeta_b200.h declares; no compute calls are made here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import _lib
    return _lib


def test_header_symbols_exported(built):
    hdr = open(os.path.join(ROOT, "include", "hymeta_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(hymeta_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    h = ctypes.CDLL(built.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(h, name), f"{name} declared in the header but not exported"
    assert declared == set(built.SIGNATURES), declared ^ set(built.SIGNATURES)


def test_version_and_status_strings(built):
    h = built.lib()
    assert h.hymeta_abi_version() == 6
    assert h.hymeta_status_string(0) == b"ok"
    for code in (-1, -2, -3, -4, -5):
        assert h.hymeta_status_string(code) not in (b"ok", b"unknown status")
    with pytest.raises(ValueError):
        built.check(-1, "x")
    with pytest.raises(built.HymetaKernelError):
        built.check(-4, "x")


def test_workspace_planner_is_pure(built):
    h = built.lib()
    # config 2: B=64, 28/4 heads, 34 blocks of 256, 7B window (4096 keys before the query)
    a = h.hymeta_swa_decode_workspace_bytes(64, 28, 4, 128, 34, 256, 4096, 0)
    b = h.hymeta_swa_decode_workspace_bytes(64, 28, 4, 128, 34, 256, 4096, 128)
    assert 0 < a < b
    assert h.hymeta_swa_decode_workspace_bytes(64, 28, 4, 64, 34, 256, 4096, 0) == 0  # D != 128
    assert h.hymeta_swa_decode_workspace_bytes(64, 28, 3, 128, 34, 256, 4096, 0) == 0  # H % Hkv


def test_ops_refuse_cpu_tensors(built):
    import torch
    from vllm_hymeta_b200 import ops
    q = torch.zeros(1, 4, 128, dtype=torch.bfloat16)
    st = torch.zeros(1, 4, 128, 128, dtype=torch.bfloat16)
    with pytest.raises(ValueError, match="no CPU path"):
        ops.gla_decode(q, q, q, q.float(), st, torch.zeros(1, dtype=torch.int32))
