"""-m gpu parity: SWA + meta-prefix varlen prefill (tcgen05) through the C-ABI vs the CPU oracle."""
import pytest
import torch

from gpu_util import DEV, assert_close
from oracle import swa as oswa

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import ops as o
    return o


def run_case(ops, q_lens, k_lens, H, Hkv, M, wl, with_q2, seed, strided=False):
    D = 128
    gen = torch.generator = torch.Generator().manual_seed(seed)
    Tq, Tk = sum(q_lens), sum(k_lens)
    cu_q = torch.tensor([0] + list(torch.tensor(q_lens).cumsum(0)), dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(k_lens).cumsum(0)), dtype=torch.int32)
    if strided:
        assert Tq == Tk
        qkv = torch.randn(Tq, (H + 2 * Hkv) * D, generator=gen).bfloat16()
        q = qkv[:, :H * D].view(Tq, H, D)
        k = qkv[:, H * D:(H + Hkv) * D].view(Tk, Hkv, D)
        v = qkv[:, (H + Hkv) * D:].view(Tk, Hkv, D)
        dqkv = qkv.to(DEV)
        dq = dqkv[:, :H * D].view(Tq, H, D)
        dk = dqkv[:, H * D:(H + Hkv) * D].view(Tk, Hkv, D)
        dv = dqkv[:, (H + Hkv) * D:].view(Tk, Hkv, D)
    else:
        q = torch.randn(Tq, H, D, generator=gen).bfloat16()
        k = torch.randn(Tk, Hkv, D, generator=gen).bfloat16()
        v = torch.randn(Tk, Hkv, D, generator=gen).bfloat16()
        dq, dk, dv = q.to(DEV), k.to(DEV), v.to(DEV)
    k2 = v2 = q2 = None
    if M:
        k2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
        v2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
        if with_q2:
            q2 = torch.randn(M, H, D, generator=gen).bfloat16()
    ref = oswa.swa_meta_prefill_varlen(q, k, v, q2, k2, v2, cu_q.tolist(), cu_k.tolist(), wl,
                                       out_dtype=torch.float32)
    dev = lambda t: None if t is None else t.to(DEV)
    out = ops.swa_meta_prefill_varlen(dq, dk, dv, dev(q2), dev(k2), dev(v2), cu_q.to(DEV),
                                      cu_k.to(DEV), max(q_lens) if q_lens else 0, wl)
    torch.cuda.synchronize()
    assert_close(out, ref, f"swa prefill {q_lens} wl={wl} M={M}")


def test_single_full_attention(ops):
    run_case(ops, [200], [200], 8, 2, 0, -1, False, 0)


def test_ragged_batch_window_meta_q2(ops):
    lens = [1, 63, 64, 65, 128, 129, 300]
    run_case(ops, lens, lens, 28, 4, 128, 40, True, 1)


def test_partial_meta_tile_and_strided_views(ops):
    run_case(ops, [700, 3], [700, 3], 28, 4, 16, 255, True, 2, strided=True)


def test_70b_window_long(ops):
    run_case(ops, [5000], [5000], 7, 1, 128, 4095, False, 3)       # TP=4 shard, W-1


def test_7b_window_long(ops):
    run_case(ops, [4500, 4200], [4500, 4200], 14, 2, 0, 4096, False, 4)   # 7B: W keys before self


def test_suffix_queries(ops):
    """Queries are the last Lq tokens of Lk keys (flash-attn bottom-right alignment)."""
    run_case(ops, [100, 17], [400, 17], 8, 2, 128, 150, False, 5)


def test_meta_queries_only(ops):
    run_case(ops, [], [], 28, 4, 128, 4095, True, 6)


def test_persistent_scheduler_many_items_and_empty_requests(ops):
    """More work items than CTAs (28 heads x 6 pair levels x 41 requests incl. the meta-query
    pseudo-request), empty requests in the middle of the batch, one request longer than the rest:
    every CTA of the persistent kernel walks several items, skips empty ones and hands TMEM / rings
    from item to item."""
    lens = [0, 257, 1, 0, 1500, 256, 255, 33, 700, 0] * 4
    run_case(ops, lens, lens, 28, 4, 128, 300, True, 7)


def test_persistent_scheduler_suffix_queries_mixed(ops):
    """Chunk-shaped work (few queries at the end of long key ranges) next to whole prompts."""
    q_lens = [16, 300, 1, 64, 5]
    k_lens = [900, 300, 2000, 64, 700]
    run_case(ops, q_lens, k_lens, 14, 2, 128, 511, False, 8)
