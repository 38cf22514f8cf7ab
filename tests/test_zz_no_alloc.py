"""-m gpu: hymeta_swa_meta_decode_paged_static -- the no-allocation variant of the SWA + meta paged
decode (split-K scratch in a static device array of the library) -- called through the C-ABI and
checked against the caller-workspace entry point on the same inputs, on two calls in a row with
different plans (so the second call meets the first one's leftovers in the scratch array).

Collected last on purpose (file name).
"""
import pytest
import torch

from gpu_util import DEV, assert_close, make_paged

pytestmark = pytest.mark.gpu

WS_LIMIT = 16 << 20        # HYMETA_SWA_DECODE_STATIC_WS_BYTES
TICKET_LIMIT = 64 << 10    # HYMETA_SWA_DECODE_STATIC_TICKET_BYTES


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import ops as o
    return o


def _inputs(lens, bs, M, H, Hkv, seed):
    D = 128
    kc, vc, tables, _, _ = make_paged(lens, Hkv, D, bs, seed=seed)
    gen = torch.Generator().manual_seed(seed + 1)
    q = torch.randn(len(lens), H, D, generator=gen).bfloat16().to(DEV)
    mk = mv = None
    if M:
        mk = torch.randn(M, Hkv, D, generator=gen).bfloat16().to(DEV)
        mv = torch.randn(M, Hkv, D, generator=gen).bfloat16().to(DEV)
    return (q, kc.to(DEV), vc.to(DEV), tables.to(DEV),
            torch.tensor(lens, dtype=torch.int32, device=DEV), mk, mv)


def _call_static(q, kc, vc, tables, lens, mk, mv, wl, out):
    """The raw C-ABI call: plain pointers, sizes and strides, no workspace."""
    from vllm_hymeta_b200 import _lib
    B, H, D = q.shape
    bs, Hkv = kc.shape[1], kc.shape[2]
    M = 0 if mk is None else mk.shape[0]
    ptr = lambda t: None if t is None else t.data_ptr()
    return _lib.lib().hymeta_swa_meta_decode_paged_static(
        ptr(q), ptr(kc), ptr(vc), ptr(tables), ptr(lens), ptr(mk), ptr(mv), ptr(out),
        B, H, Hkv, D, M, bs, kc.shape[0], tables.shape[1], tables.stride(0), wl, float(D ** -0.5),
        q.stride(0), q.stride(1), out.stride(0), out.stride(1),
        kc.stride(0), kc.stride(1), kc.stride(2), vc.stride(0), vc.stride(1), vc.stride(2),
        mk.stride(0) if M else 0, mk.stride(1) if M else 0,
        mv.stride(0) if M else 0, mv.stride(1) if M else 0,
        torch.cuda.current_stream().cuda_stream)


# (lens, block_size, window_left, meta tokens): few long requests -> many splits + tickets;
# a batch with ragged lengths, an empty (padded) request and meta keys -> another plan
CALLS = [([5000, 4097, 1], 256, 4096, 0),
         ([700, 0, 31, 2600, 129, 64, 9000], 64, 1023, 128)]


def test_static_scratch_decode_matches_workspace_entry_two_calls_in_a_row(ops):
    H, Hkv = 28, 4
    for rnd in range(2):                      # the whole sequence twice: state left behind is harmless
        for i, (lens, bs, wl, M) in enumerate(CALLS):
            q, kc, vc, tables, sl, mk, mv = _inputs(lens, bs, M, H, Hkv, seed=11 + i)
            ref = ops.swa_meta_decode_paged(q, kc, vc, tables, sl, mk, mv, wl)
            out = torch.full_like(ref, float("nan"))
            torch.cuda.synchronize()
            before = torch.cuda.memory_allocated()
            rc = _call_static(q, kc, vc, tables, sl, mk, mv, wl, out)
            assert rc == 0, f"hymeta_swa_meta_decode_paged_static returned {rc}"
            torch.cuda.synchronize()
            assert torch.cuda.memory_allocated() == before
            assert_close(out, ref, f"static-scratch decode, round {rnd} call {i}")
            # same kernel, same plan, same summation order: the rows are the same bits
            assert torch.equal(out, ref)


def test_static_scratch_decode_rejects_plans_beyond_the_array(ops):
    """Sizes whose partials or tickets do not fit the static array return HYMETA_ERR_WORKSPACE (-3)
    before anything is launched (the pointers are never dereferenced)."""
    from vllm_hymeta_b200 import _lib
    h = _lib.lib()
    H, Hkv, D, bs, wl = 28, 4, 128, 256, 4096
    lens, bsz, _, M = CALLS[0]
    q, kc, vc, tables, sl, mk, mv = _inputs(lens, bsz, M, H, Hkv, seed=3)
    out = torch.zeros_like(q)

    def rc_for(B, max_blocks):
        return h.hymeta_swa_meta_decode_paged_static(
            q.data_ptr(), kc.data_ptr(), vc.data_ptr(), tables.data_ptr(), sl.data_ptr(), None, None,
            out.data_ptr(), B, H, Hkv, D, 0, bs, kc.shape[0], max_blocks, max_blocks, wl, float(D ** -0.5),
            q.stride(0), q.stride(1), out.stride(0), out.stride(1),
            kc.stride(0), kc.stride(1), kc.stride(2), vc.stride(0), vc.stride(1), vc.stride(2),
            0, 0, 0, 0, torch.cuda.current_stream().cuda_stream)

    B_big = 2000                                  # 2000 x 28 rows x 516 B of partials > 16 MiB
    assert h.hymeta_swa_decode_workspace_bytes(B_big, H, Hkv, D, 32, bs, wl, 0) > WS_LIMIT
    assert rc_for(B_big, 32) == -3
    B_tickets = 20000                             # 20000 x 4 tickets x 4 B > 64 KiB
    assert h.hymeta_swa_decode_ticket_bytes(B_tickets, H, Hkv) > TICKET_LIMIT
    assert rc_for(B_tickets, 1) == -3
    torch.cuda.synchronize()
    # and the array still serves the next in-range call
    ref = ops.swa_meta_decode_paged(q, kc, vc, tables, sl, mk, mv, wl)
    got = torch.empty_like(ref)
    assert _call_static(q, kc, vc, tables, sl, mk, mv, wl, got) == 0
    torch.cuda.synchronize()
    assert torch.equal(got, ref)
