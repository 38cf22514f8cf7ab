"""-m gpu parity: SWA + meta-prefix prefill with keys read from the PAGED cache through the block
table -- the prefix-cache / chunked-prefill branch (reference meta_attention.py:886-919) and the
multi-token decode branch (:924-953) -- through the C-ABI vs ``oracle.swa.swa_meta_prefill_paged``.
Unowned cache rows (tail of the last page, pages outside the block table) hold NaN: a kernel that
lets them reach the MMA poisons its output."""
import pytest
import torch

from gpu_util import DEV, assert_close
from oracle import swa as oswa

pytestmark = pytest.mark.gpu
D = 128


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import ops as o
    return o


def paged_cache(seq_lens, Hkv, bs, seed, extra_blocks=5, max_blocks=None, fill=float("nan")):
    """Caches [nb, bs, Hkv, D] filled with ``fill``; request n owns ceil(L/bs) pages taken from a
    seeded permutation and holds N(0,1) keys / values at its first L logical positions."""
    gen = torch.Generator().manual_seed(seed)
    nblk = [(L + bs - 1) // bs for L in seq_lens]
    nb = sum(nblk) + extra_blocks
    kc = torch.full((nb, bs, Hkv, D), fill, dtype=torch.bfloat16)
    vc = torch.full((nb, bs, Hkv, D), fill, dtype=torch.bfloat16)
    perm = torch.randperm(nb, generator=gen)
    mb = max_blocks or max(max(nblk), 1)
    tables = torch.zeros(len(seq_lens), mb, dtype=torch.int32)
    at = 0
    for n, L in enumerate(seq_lens):
        row = perm[at:at + nblk[n]]
        at += nblk[n]
        tables[n, :nblk[n]] = row.to(torch.int32)
        pos = torch.arange(L)
        kc[row[pos // bs], pos % bs] = torch.randn(L, Hkv, D, generator=gen).bfloat16()
        vc[row[pos // bs], pos % bs] = torch.randn(L, Hkv, D, generator=gen).bfloat16()
    return kc, vc, tables


def run_case(ops, q_lens, seq_lens, H, Hkv, M, wl, bs, seed, with_q2=False, check_permuted=True):
    gen = torch.Generator().manual_seed(seed)
    kc, vc, tables = paged_cache(seq_lens, Hkv, bs, seed + 1000)
    Tq = sum(q_lens)
    q = torch.randn(Tq, H, D, generator=gen).bfloat16()
    k2 = v2 = q2 = None
    if M:
        k2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
        v2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
        if with_q2:
            q2 = torch.randn(M, H, D, generator=gen).bfloat16()
    cu_q = torch.tensor([0] + torch.tensor(q_lens).cumsum(0).tolist(), dtype=torch.int32)
    cu_k = torch.tensor([0] + torch.tensor(seq_lens).cumsum(0).tolist(), dtype=torch.int32)
    ref = oswa.swa_meta_prefill_paged(q, kc, vc, tables.tolist(), cu_q.tolist(), seq_lens, k2, v2, wl,
                                      out_dtype=torch.float32)
    if q2 is not None:
        om, _ = oswa._attend(q2, oswa._rep(k2, H), oswa._rep(v2, H),
                             torch.ones(M, M, dtype=torch.bool), D ** -0.5)   # bidirectional (:267)
        ref = torch.cat([om, ref], 0)
    dev = lambda t: None if t is None else t.to(DEV)
    out = ops.swa_meta_prefill_varlen(dev(q), dev(kc), dev(vc), dev(q2), dev(k2), dev(v2), dev(cu_q),
                                      dev(cu_k), max(q_lens), wl, block_tables=dev(tables))
    torch.cuda.synchronize()
    assert_close(out, ref, f"paged swa prefill q={q_lens} L={seq_lens} wl={wl} M={M} bs={bs}")
    if check_permuted:
        # the same logical cache behind a different physical page order: bit-identical output
        nb = kc.shape[0]
        perm = torch.randperm(nb, generator=gen)
        inv = torch.empty_like(perm)
        inv[perm] = torch.arange(nb)
        kc2, vc2 = kc[perm], vc[perm]                 # new block j holds old block perm[j]
        tables2 = inv[tables.long()].to(torch.int32)  # old block id -> its new position
        out2 = ops.swa_meta_prefill_varlen(dev(q), dev(kc2), dev(vc2), dev(q2), dev(k2), dev(v2),
                                           dev(cu_q), dev(cu_k), max(q_lens), wl,
                                           block_tables=dev(tables2))
        torch.cuda.synchronize()
        assert torch.equal(out.cpu(), out2.cpu()), "output depends on the physical page order"
    return out


@pytest.mark.parametrize("bs", [16, 256])
def test_ragged_prefix_hit_window_meta(ops, bs):
    """Ragged context_len > 0 (queries are the last tokens), context 0 and a 1-token chunk."""
    run_case(ops, [100, 17, 260, 1, 64], [400, 17, 300, 129, 64 + 512], 14, 2, 128, 150, bs, 1)


@pytest.mark.parametrize("bs", [16, 32, 64, 128, 256])
def test_block_sizes_full_attention_no_meta(ops, bs):
    run_case(ops, [130, 40], [700, 41], 8, 2, 0, -1, bs, 2)


@pytest.mark.parametrize("wl,M", [(4095, 128), (4096, 0), (-1, 128)])
def test_chunked_prefill_long_context(ops, wl, M):
    """Second chunk of a long prompt (70B: W-1 = 4095 with meta; 7B: W = 4096 without; a
    full-attention layer): the window start falls inside a page and inside a 128-key tile."""
    run_case(ops, [300, 129], [4500 + 300, 5000], 7, 1, M, wl, 256, 3)


def test_meta_queries_with_paged_keys(ops):
    run_case(ops, [70, 200], [70, 333], 14, 2, 128, 63, 16, 4, with_q2=True)


def test_multi_token_decode_shape(ops):
    """max_decode_query_len > 1 (speculative / multi-token decode, :924-953): a few queries per
    request at the end of a long paged context."""
    run_case(ops, [4, 3, 1, 4], [1000, 259, 16, 4100], 28, 4, 128, 4095, 256, 5)


def test_matches_packed_kernel_bitwise(ops):
    """Same keys packed and paged: the two producers feed the same tiles, so the outputs are equal
    bit for bit (context 0, so the packed form applies)."""
    lens = [300, 129, 1000]
    H, Hkv, M, wl, bs = 14, 2, 128, 200, 64
    gen = torch.Generator().manual_seed(6)
    kc, vc, tables = paged_cache(lens, Hkv, bs, 77)
    q = torch.randn(sum(lens), H, D, generator=gen).bfloat16()
    k2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
    v2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32)
    kp = torch.cat([oswa.gather_paged(kc, tables[n].tolist(), 0, L) for n, L in enumerate(lens)])
    vp = torch.cat([oswa.gather_paged(vc, tables[n].tolist(), 0, L) for n, L in enumerate(lens)])
    a = ops.swa_meta_prefill_varlen(q.to(DEV), kc.to(DEV), vc.to(DEV), None, k2.to(DEV), v2.to(DEV),
                                    cu.to(DEV), cu.to(DEV), max(lens), wl, block_tables=tables.to(DEV))
    b = ops.swa_meta_prefill_varlen(q.to(DEV), kp.to(DEV), vp.to(DEV), None, k2.to(DEV), v2.to(DEV),
                                    cu.to(DEV), cu.to(DEV), max(lens), wl)
    torch.cuda.synchronize()
    assert torch.equal(a.cpu(), b.cpu())


def test_bad_arguments(ops):
    q = torch.zeros(4, 2, D, dtype=torch.bfloat16, device=DEV)
    cu = torch.tensor([0, 4], dtype=torch.int32, device=DEV)
    bt = torch.zeros(1, 1, dtype=torch.int32, device=DEV)
    cache = torch.zeros(2, 24, 1, D, dtype=torch.bfloat16, device=DEV)     # block_size 24: not % 16
    with pytest.raises(ValueError):
        ops.swa_meta_prefill_varlen(q, cache, cache, None, None, None, cu, cu, 4, -1, block_tables=bt)
    with pytest.raises(ValueError):
        ops.swa_meta_prefill_varlen(q, cache[0], cache[0], None, None, None, cu, cu, 4, -1, block_tables=bt)
