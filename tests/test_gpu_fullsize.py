"""-m gpu parity at BASELINE.json's FULL sizes (configs 2-5), where the CPU oracle cannot run the
whole case: inputs are generated on the device, the kernels run at full size, and the results are
checked (1) on SAMPLED rows against the fp32 oracle evaluated only for those rows, and (2) through
size-independent properties of the domain -- prefix consistency of the recurrence, window locality,
prefill/decode agreement, bit-exact block-table indirection (a permuted physical layout must give
bit-identical outputs)."""
import pytest
import torch

from oracle import gla as ogla
from oracle import swa as oswa
from gpu_util import DEV, assert_close

pytestmark = pytest.mark.gpu
D = 128


@pytest.fixture(scope="module")
def ops():
    from vllm_hymeta_b200 import ops as o
    return o


def _i32(x):
    return torch.tensor(x, dtype=torch.int32, device=DEV)


# ------------------------------------------------------------------ config 3: 70B shard, 32K + meta
@pytest.mark.parametrize("W", [4095, -1])
def test_swa_meta_prefill_32k_sampled_rows_and_window_locality(ops, W):
    T, H, Hkv, M = 32768, 7, 1, 128                       # TP=4 shard of the 28/4-head layer
    gen = torch.Generator(device=DEV).manual_seed(3)
    q, k, v = (torch.randn(T, h, D, generator=gen, device=DEV).bfloat16() for h in (H, Hkv, Hkv))
    k2, v2 = (torch.randn(M, Hkv, D, generator=gen, device=DEV).bfloat16() for _ in range(2))
    cu = _i32([0, T])
    out = ops.swa_meta_prefill_varlen(q, k, v, None, k2, v2, cu, cu, T, W)
    torch.cuda.synchronize()
    rows = [0, 1, 127, 128, 4094, 4095, 4096, 4097, 12345, 20000, 32767 - 4096, 32766, 32767]
    for i in rows:
        lo = 0 if W < 0 else max(0, i - W)
        ref, _ = oswa.swa_meta_dense(q[i:i + 1].cpu(), k[lo:i + 1].cpu(), v[lo:i + 1].cpu(), None,
                                     k2.cpu(), v2.cpu(), -1)
        assert_close(out[i:i + 1], ref, f"32K row {i} W={W}")
    if W >= 0:
        # window locality: a query at i >= s + W never sees keys before s, so running only the
        # suffix [s, T) must reproduce those rows (same math, different tiling -> tolerance)
        s = 16384
        cu2 = _i32([0, T - s])
        out2 = ops.swa_meta_prefill_varlen(q[s:], k[s:], v[s:], None, k2, v2, cu2, cu2, T - s, W)
        torch.cuda.synchronize()
        assert_close(out2[W:], out[s + W:].float(), "window locality (suffix run)",
                     maxabs_frac=2.0 ** -7, rel_rms=4e-3)


# ------------------------------------------------------------------ config 2/4: paged decode
def _ragged_paged(lens, Hkv, bs, gen):
    """Device-side paged cache with a random physical permutation; returns caches, tables, and the
    logical->physical map used to read rows back."""
    nblk = [(L + bs - 1) // bs for L in lens]
    nb = sum(nblk) + 5
    kc = torch.randn(nb, bs, Hkv, D, generator=gen, device=DEV).bfloat16()
    vc = torch.randn(nb, bs, Hkv, D, generator=gen, device=DEV).bfloat16()
    perm = torch.randperm(nb, generator=gen, device=DEV).to(torch.int32)
    tables = torch.zeros(len(lens), max(nblk), dtype=torch.int32, device=DEV)
    off = 0
    for i, n in enumerate(nblk):
        tables[i, :n] = perm[off:off + n]
        off += n
    return kc, vc, tables


def test_swa_decode_config4_1000_ragged_seqs(ops):
    """1000 concurrent sequences, lengths U[1K, 9K], 256-token blocks (config 4): sampled requests
    vs the oracle over their gathered window, and a re-permuted physical layout must give
    BIT-IDENTICAL outputs (block-table indirection is exact index work)."""
    B, H, Hkv, bs, W = 1000, 28, 4, 256, 4096
    gen = torch.Generator(device=DEV).manual_seed(4)
    lens = torch.randint(1024, 9217, (B,), generator=gen, device=DEV).tolist()
    kc, vc, tables = _ragged_paged(lens, Hkv, bs, gen)
    q = torch.randn(B, H, D, generator=gen, device=DEV).bfloat16()
    sl = _i32(lens)
    out = ops.swa_meta_decode_paged(q, kc, vc, tables, sl, None, None, W)
    torch.cuda.synchronize()
    for b in [0, 1, 17, 499, 998, 999, int(torch.tensor(lens).argmax()), int(torch.tensor(lens).argmin())]:
        L = lens[b]
        lo = max(0, L - 1 - W)
        pos = torch.arange(lo, L, device=DEV)
        blk = tables[b, (pos // bs).long()].long()
        kk, vv = kc[blk, (pos % bs).long()].cpu(), vc[blk, (pos % bs).long()].cpu()
        ref, _ = oswa.swa_meta_dense(q[b:b + 1].cpu(), kk, vv, None, None, None, -1)
        assert_close(out[b:b + 1], ref, f"config-4 request {b} (len {L})")
    # same logical content behind a different physical permutation
    nb = kc.shape[0]
    perm = torch.randperm(nb, generator=gen, device=DEV)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(nb, device=DEV)
    kc2, vc2 = kc[perm], vc[perm]                         # new block j holds old block perm[j]
    tables2 = inv[tables.long()].to(torch.int32)
    out2 = ops.swa_meta_decode_paged(q, kc2, vc2, tables2, sl, None, None, W)
    torch.cuda.synchronize()
    assert torch.equal(out, out2), "block-table indirection must be bit-exact"


def test_swa_prefill_then_decode_agree_at_8k(ops):
    """Config 2 shape per request (T = 8192, window 4096): the decode kernel over the paged cache at
    ctx 8192 must agree with the last row of the prefill kernel over the same keys."""
    T, H, Hkv, bs, W = 8192, 28, 4, 256, 4096
    gen = torch.Generator(device=DEV).manual_seed(5)
    q, k, v = (torch.randn(T, h, D, generator=gen, device=DEV).bfloat16() for h in (H, Hkv, Hkv))
    cu = _i32([0, T])
    o_pf = ops.swa_meta_prefill_varlen(q, k, v, None, None, None, cu, cu, T, W)
    nb = T // bs
    perm = torch.randperm(nb + 3, generator=gen, device=DEV)[:nb]
    cache = torch.zeros(2, nb + 3, bs, Hkv, D, dtype=torch.bfloat16, device=DEV)
    slots = (perm[:, None] * bs + torch.arange(bs, device=DEV)[None]).reshape(-1).to(torch.int64)
    ops.kv_write_paged(k, v, cache[0], cache[1], slots)
    o_dec = ops.swa_meta_decode_paged(q[T - 1:T], cache[0], cache[1], perm[None].to(torch.int32),
                                      _i32([T]), None, None, W)
    torch.cuda.synchronize()
    assert_close(o_dec, o_pf[T - 1:T].float(), "decode vs last prefill row", rel_rms=4e-3)


# ------------------------------------------------------------------ config 5: GLA long-context scan
@pytest.mark.parametrize("T", [32768, 131072])
def test_gla_prefill_long_context_prefix_and_continuation(ops, T):
    """One 7B kv-head shard (H=7, Hkv=1) over T tokens: (1) the first 2048 outputs match the fp32
    oracle recurrence; (2) prefix property -- splitting the sequence into two runs chained through
    the slot state reproduces the single run; (3) the segment-parallel form (summaries -> scan ->
    outputs) agrees with the plain chain."""
    H, Hkv = 7, 1
    gen = torch.Generator(device=DEV).manual_seed(T)
    q = torch.randn(T, H, D, generator=gen, device=DEV).bfloat16().relu()
    k = torch.randn(T, Hkv, D, generator=gen, device=DEV).bfloat16().relu()
    v = torch.randn(T, Hkv, D, generator=gen, device=DEV).bfloat16()
    g = (torch.nn.functional.logsigmoid(torch.randn(T, Hkv, D, generator=gen, device=DEV)) / 16).bfloat16()
    st = lambda: torch.zeros(1, Hkv, D, D, dtype=torch.bfloat16, device=DEV)
    s_chain, s_seg, s_split = st(), st(), st()
    one = dict(max_seqlen=T)
    o_chain = ops.gla_prefill_varlen(q, k, v, g, s_chain, _i32([0, T]), _i32([0]), _i32([0]), seg_len=0, **one)
    o_seg = ops.gla_prefill_varlen(q, k, v, g, s_seg, _i32([0, T]), _i32([0]), _i32([0]), seg_len=1024, **one)
    torch.cuda.synchronize()
    n = 2048
    o_ref, _ = ogla.gla_chunked(q[:n].cpu(), k[:n].cpu(), v[:n].cpu(), g[:n].cpu(), chunk=64)
    assert_close(o_chain[:n], o_ref, f"GLA T={T}: first {n} rows vs oracle")
    assert_close(o_seg, o_chain.float(), f"GLA T={T}: segment-parallel vs chain", rel_rms=4e-3)
    assert_close(s_seg, s_chain.float(), f"GLA T={T}: final state, segment-parallel vs chain", rel_rms=4e-3)
    cut = T // 2 + 16 * 5                                  # a step boundary (C = 16) past the middle
    a = ops.gla_prefill_varlen(q[:cut], k[:cut], v[:cut], g[:cut], s_split, _i32([0, cut]), _i32([0]),
                               _i32([0]), seg_len=0, max_seqlen=cut)
    b = ops.gla_prefill_varlen(q[cut:], k[cut:], v[cut:], g[cut:], s_split, _i32([0, T - cut]), _i32([0]),
                               _i32([1]), seg_len=0, max_seqlen=T - cut)        # init from the slot
    torch.cuda.synchronize()
    assert torch.equal(a, o_chain[:cut]), "prefix rows must be bit-identical"
    # the continuation restarts from the bf16-rounded slot state: within bf16 state rounding
    assert_close(b, o_chain[cut:].float(), f"GLA T={T}: continuation through the slot state", rel_rms=4e-3)
    assert_close(s_split, s_chain.float(), f"GLA T={T}: final state after the split run", rel_rms=4e-3)


# ------------------------------------------------------------------ config 2: the full bench batch
def test_config2_full_batch_prefill_64x8192_sampled_rows(ops):
    """The bench.py launch shapes themselves: 64 packed sequences x 8192 tokens (524 288 rows) in ONE
    launch of each prefill kernel at the 7B geometry (28/4 heads, window 4096 keys before self,
    no meta) -- the ``blockIdx.z = n`` x ``n_qpairs`` grid and the 256-chain GLA grid at benchmark
    size.  Sampled rows of sampled sequences against the fp32 oracle evaluated for those rows, the
    GLA final states of two sequences against the oracle recurrence, and every sequence's rows
    checked finite."""
    N, L, H, Hkv, W = 64, 8192, 28, 4, 4096
    T = N * L
    gen = torch.Generator(device=DEV).manual_seed(64)
    cu = _i32([i * L for i in range(N + 1)])
    # ---- SWA prefill
    q, k, v = (torch.randn(T, h, D, generator=gen, device=DEV).bfloat16() for h in (H, Hkv, Hkv))
    out = ops.swa_meta_prefill_varlen(q, k, v, None, None, None, cu, cu, L, W)
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all()
    for n in (0, 1, 31, 63):
        for i in (0, 127, 128, 255, 256, 4095, 4096, 4097, 6000, 8191):
            lo = max(0, i - W)
            r = n * L
            ref, _ = oswa.swa_meta_dense(q[r + i:r + i + 1].cpu(), k[r + lo:r + i + 1].cpu(),
                                         v[r + lo:r + i + 1].cpu(), None, None, None, -1)
            assert_close(out[r + i:r + i + 1], ref, f"config-2 SWA prefill seq {n} row {i}")
    del out
    # ---- GLA prefill (7B gate: relu q/k, g = logsigmoid/16), slot states written in-kernel
    qg = torch.relu(q)
    kg = torch.relu(k)
    g = (torch.nn.functional.logsigmoid(torch.randn(T, Hkv, D, generator=gen, device=DEV)) / 16).bfloat16()
    state = torch.zeros(N, Hkv, D, D, dtype=torch.bfloat16, device=DEV)
    slots = torch.randperm(N, generator=gen, device=DEV).to(torch.int32)
    o = ops.gla_prefill_varlen(qg, kg, v, g, state, cu, slots, _i32([0] * N), max_seqlen=L)
    torch.cuda.synchronize()
    assert torch.isfinite(o.float()).all()
    for n in (0, 63):
        r = slice(n * L, (n + 1) * L)
        ro, rS = ogla.gla_chunked(qg[r].cpu(), kg[r].cpu(), v[r].cpu(), g[r].cpu())
        rows = [0, 1, 15, 16, 63, 64, 1000, 4096, 8191]
        assert_close(o[r][rows], ro[rows], f"config-2 GLA prefill seq {n} sampled rows")
        assert_close(state[int(slots[n])], rS[:: H // Hkv], f"config-2 GLA final state seq {n}")
    # the launch above ran the balanced work lists (256 chains > 148 SMs, chains cut across SMs); one
    # chain per CTA must give the same bits for every row and every state
    state1 = torch.zeros_like(state)
    o1 = ops.gla_prefill_varlen(qg, kg, v, g, state1, cu, slots, _i32([0] * N), max_seqlen=L, seg_len=0)
    torch.cuda.synchronize()
    assert torch.equal(o1, o) and torch.equal(state1, state)
