"""-m gpu: the a5 device kernels -- ``hymeta_state_zero_slots`` (reference ``_clear_prefill_cache``,
modeling_hymeta.py:1202-1227 / modeling_gla_swa.py:391-414: ``cache[:, slots] = 0``) and
``hymeta_state_copy_slot`` (``_copy_cache``, hymeta_cache.py:46-50: ``cache[:, to] = cache[:, from]``)
-- through the C-ABI, bit-exact against the torch indexing they replace, plus the two cache
managers driving them on device tensors."""
import pytest
import torch

from gpu_util import DEV

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import ops as o
    return o


def _state(L, S, Hs, D, seed, pad_layers=0, pad_slots=0):
    """[L, S, Hs, D, D] bf16 view with NON-compact layer / slot strides (a slice of a larger
    allocation), and the full parent so bytes outside the view can be checked too."""
    gen = torch.Generator().manual_seed(seed)
    parent = torch.randn(L + pad_layers, S + pad_slots, Hs, D, D, generator=gen).bfloat16().to(DEV)
    return parent, parent[:L, :S]


@pytest.mark.parametrize("shape", [(1, 4, 1, 128), (14, 9, 4, 128), (3, 70, 28, 128)])
@pytest.mark.parametrize("strided", [False, True])
def test_state_zero_slots_bit_exact(ops, shape, strided):
    L, S, Hs, D = shape
    parent, state = _state(L, S, Hs, D, 1, 2 if strided else 0, 3 if strided else 0)
    ref_parent = parent.clone()
    gen = torch.Generator().manual_seed(2)
    pick = torch.randperm(S, generator=gen)[:max(1, S // 2)].tolist()
    slots = pick + [-1, -1]                       # PAD entries are skipped
    ref_parent[:L, pick] = 0
    ops.state_zero_slots(state, torch.tensor(slots, dtype=torch.int32, device=DEV))
    torch.cuda.synchronize()
    assert torch.equal(parent.view(torch.int16), ref_parent.view(torch.int16))
    # empty list and all-PAD list are no-ops
    ops.state_zero_slots(state, torch.tensor([], dtype=torch.int32, device=DEV))
    ops.state_zero_slots(state, torch.tensor([-1], dtype=torch.int32, device=DEV))
    torch.cuda.synchronize()
    assert torch.equal(parent.view(torch.int16), ref_parent.view(torch.int16))


@pytest.mark.parametrize("shape", [(1, 4, 1, 128), (14, 9, 4, 128), (28, 5, 28, 128)])
@pytest.mark.parametrize("strided", [False, True])
def test_state_copy_slot_bit_exact(ops, shape, strided):
    L, S, Hs, D = shape
    parent, state = _state(L, S, Hs, D, 3, 1 if strided else 0, 2 if strided else 0)
    ref_parent = parent.clone()
    src, dst = 1, S - 1
    ref_parent[:L, dst] = ref_parent[:L, src]
    ops.state_copy_slot(state, src, dst)
    ops.state_copy_slot(state, 2, 2)              # self copy: no-op
    torch.cuda.synchronize()
    assert torch.equal(parent.view(torch.int16), ref_parent.view(torch.int16))


def test_state_ops_reject_bad_arguments(ops):
    _, state = _state(2, 3, 2, 128, 4)
    with pytest.raises(ValueError):
        ops.state_zero_slots(state, torch.tensor([0], dtype=torch.int64, device=DEV))
    with pytest.raises(ValueError):
        ops.state_zero_slots(state.cpu(), torch.tensor([0], dtype=torch.int32))
    with pytest.raises(ValueError):
        ops.state_copy_slot(state, -1, 1)


def test_cache_managers_on_device(ops):
    """HymetaCacheManager / GLACacheManager (hymeta_cache.py:24-63, gla_cache.py:17-32) over device
    tensors: admit, fork (sibling sequence copies the parent's state in every layer), finish,
    recycle + clear."""
    from vllm_hymeta_b200.model_for_7B.gla_cache import GLACacheManager
    from vllm_hymeta_b200.models.hymeta_cache import HymetaCacheManager
    L, S, Hs, D, M, Hkv = 3, 4, 2, 128, 16, 1
    m = HymetaCacheManager(torch.bfloat16, (L, S, Hs, D, D), (L, Hs, D, D), (L, 2, M, Hkv, D), device=DEV)
    state, _, _, idx = m.current_run_tensors(request_ids_to_seq_ids={"a": [0]}, finished_requests_ids=[])
    assert state.is_cuda and state.abs().max().item() == 0 and idx.dtype == torch.int32
    a = int(idx[0])
    gen = torch.Generator().manual_seed(6)
    state[:, a] = torch.randn(L, Hs, D, D, generator=gen).bfloat16().to(DEV)
    _, _, _, idx2 = m.current_run_tensors(request_ids_to_seq_ids={"a": [0, 5]}, finished_requests_ids=[])
    b = int(idx2[1])
    torch.cuda.synchronize()
    assert b != a and torch.equal(state[:, b], state[:, a])             # forked on device
    _, _, _, idx3 = m.current_run_tensors(request_ids_to_seq_ids={"c": [9]}, finished_requests_ids=["a"])
    c = int(idx3[0])
    assert c in (a, b)
    m.clear_slots([c])
    torch.cuda.synchronize()
    assert state[:, c].abs().max().item() == 0
    g = GLACacheManager(torch.bfloat16, (L, S, Hs, D, D), device=DEV)
    cache, gi = g.current_run_tensors(request_ids_to_seq_ids={"x": [0]}, finished_requests_ids=[])
    cache[:, int(gi[0])] = 2.0
    g._copy_cache(int(gi[0]), (int(gi[0]) + 1) % S)                      # copies SLOTS
    torch.cuda.synchronize()
    assert cache[:, (int(gi[0]) + 1) % S].min().item() == 2
