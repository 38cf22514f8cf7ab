"""-m gpu parity at the MODULE boundary: gate pre-processing kernels, the GLA cores
(70B ``HLinearAttention`` / 7B ``GatedLinearAttention``) on mixed [meta | prefill | decode]
batches with slot states, and ``FlashAttention`` -> ``MetaAttentionImpl.forward`` over the paged
cache -- each against the CPU oracle composed the way the reference composes its torch path."""
from types import SimpleNamespace

import pytest
import torch

from gpu_util import DEV, assert_close
from oracle import gla as ogla
from oracle import swa as oswa

pytestmark = pytest.mark.gpu
D = 128


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import ops as o
    return o


def _qkv(T, H, Hkv, seed):
    gen = torch.Generator().manual_seed(seed)
    return torch.randn(T, (H + 2 * Hkv) * D, generator=gen).bfloat16()


# ------------------------------------------------------------------------------------- a4
@pytest.mark.parametrize("T", [1, 37, 1000])
def test_gate_70b_kernel(ops, T):
    H, Hkv = 14, 2
    qkv = _qkv(T, H, Hkv, 3)
    q = qkv[:, :H * D].view(T, H, D)
    k = qkv[:, H * D:(H + Hkv) * D].view(T, Hkv, D)
    v = qkv[:, (H + Hkv) * D:].view(T, Hkv, D)
    dq = qkv.to(DEV)
    dqv = dq[:, :H * D].view(T, H, D)
    dkv = dq[:, H * D:(H + Hkv) * D].view(T, Hkv, D)
    # no lower bound: per kv-head outputs
    rq, rk, rg = ogla.gate_70b(q, k, v)
    oq, ok, og = ops.gla_gate_70b(dqv, dkv)
    assert torch.equal(oq.cpu(), rq) and og.dtype == torch.float32 and ok.shape == (T, Hkv, D)
    assert (ok.float().cpu() - rk.float()).abs().max() <= 2 ** -8      # 1 bf16 ulp at k' <= 0.95
    assert torch.allclose(og.cpu(), rg, atol=2e-6, rtol=2e-6)
    # lower bound (per q-head, rank-local slice): k', g per q-head, raw k broadcast in-kernel
    gen = torch.Generator().manual_seed(5)
    lb_full = torch.rand(2 * H * D, generator=gen) * 0.5               # 2 ranks' worth of heads
    rq, rk, rg = ogla.gate_70b(q, k.repeat_interleave(H // Hkv, 1), v, lower_bound=lb_full,
                               layer_idx=3, num_heads=2 * H, head_start=H)
    lb_local = lb_full.view(2 * H, D)[H:].contiguous().to(DEV)
    oq, ok, og = ops.gla_gate_70b(dqv, dkv, lb_local)
    assert ok.shape == (T, H, D) and og.shape == (T, H, D)
    assert (ok.float().cpu() - rk.float()).abs().max() <= 2 ** -8
    assert torch.allclose(og.cpu(), rg, atol=2e-6, rtol=2e-6)


@pytest.mark.parametrize("T", [1, 37, 1000])
def test_gate_7b_kernel(ops, T):
    H, Hkv = 28, 4
    qkv = _qkv(T, H, Hkv, 7)
    gk = torch.randn(T, Hkv * D, generator=torch.Generator().manual_seed(8)).bfloat16() * 3
    q = qkv[:, :H * D].view(T, H, D)
    k = qkv[:, H * D:(H + Hkv) * D].view(T, Hkv, D)
    rq, rk, rg = ogla.gate_7b(q, k, gk.view(T, Hkv, D))
    dq = qkv.to(DEV)
    oq, ok, og = ops.gla_gate_7b(dq[:, :H * D].view(T, H, D), dq[:, H * D:(H + Hkv) * D].view(T, Hkv, D),
                                 gk.to(DEV).view(T, Hkv, D))
    assert torch.equal(oq.cpu(), rq) and torch.equal(ok.cpu(), rk)
    assert og.dtype == torch.bfloat16
    # bf16 logsigmoid then /16: allow one bf16 ulp where the fp32 value sits on a rounding boundary
    d = (og.float().cpu() - rg.float()).abs()
    assert (d <= rg.float().abs() * 2 ** -7 + 1e-9).all()
    assert (d == 0).float().mean() > 0.99


# ------------------------------------------------------------------------------------- a3 + a2 (70B core)
def _md(num_prefills, prefill_lens, num_decodes, ctx):
    qsl = [0]
    for L in prefill_lens:
        qsl.append(qsl[-1] + L)
    for _ in range(num_decodes):
        qsl.append(qsl[-1] + 1)
    return SimpleNamespace(
        num_prefills=num_prefills, num_prefill_tokens=sum(prefill_lens), num_decode_tokens=num_decodes,
        query_start_loc=torch.tensor(qsl, dtype=torch.int32, device=DEV),
        context_lens_tensor=torch.tensor(ctx, dtype=torch.int32, device=DEV),
        max_query_len=max(prefill_lens + [1]))


@pytest.mark.parametrize("use_lb", [False, True])
def test_hlinear_attention_mixed_batch_with_meta_warmup(ops, use_lb):
    from vllm_hymeta_b200.models.hymeta_cache import HymetaCacheParams
    from vllm_hymeta_b200.models.linear_attention import HLinearAttention
    Htot, Hkvtot, tp, rank, M = 28, 4, 2, 1, 128
    H, Hkv = Htot // tp, Hkvtot // tp
    lens, ctx, n_dec = [70, 33], [0, 48], 3
    T = M + sum(lens) + n_dec
    qkv = _qkv(T, H, Hkv, 11)
    Hs = H if use_lb else Hkv
    gen = torch.Generator().manual_seed(12)
    nslots = 8
    state = (torch.randn(nslots, Hs, D, D, generator=gen) * 0.2).bfloat16()
    slots = [5, 2, 7, 0, 3]                       # [prefill0, prefill1, dec0, dec1, dec2]
    lb_full = torch.rand(Htot * D, generator=gen) * 0.5 if use_lb else None
    layer_idx = 3

    # ---- oracle, composed like the reference (modeling_hymeta.py:441-659)
    q = qkv[:, :H * D].view(T, H, D)
    k = qkv[:, H * D:(H + Hkv) * D].view(T, Hkv, D)
    v = qkv[:, (H + Hkv) * D:].view(T, Hkv, D)
    if use_lb:
        rq, rk, rg = ogla.gate_70b(q, k.repeat_interleave(H // Hkv, 1), v, lower_bound=lb_full,
                                   layer_idx=layer_idx, num_heads=Htot, head_start=rank * H)
        rv = v.repeat_interleave(H // Hkv, 1)
    else:
        rq, rk, rg = ogla.gate_70b(q, k, v)
        rv = v
    ref_state = state.clone()
    o_meta, S_meta = ogla.gla_recurrent(rq[:M], rk[:M], rv[:M], rg[:M])
    S_meta = S_meta[:: S_meta.shape[0] // Hs].bfloat16()               # meta_cache.copy_(state)
    cu = [0, lens[0], lens[0] + lens[1]]
    modes = [ogla.INIT_META if c == 0 else ogla.INIT_SLOT for c in ctx]
    sl = slice(M, M + sum(lens))
    o_pf = ogla.gla_prefill_varlen(rq[sl], rk[sl], rv[sl], rg[sl], ref_state, cu, slots[:2], modes,
                                   S_meta, out_dtype=torch.float32)
    ds = slice(M + sum(lens), T)
    o_dec = ogla.gla_decode_step(rq[ds], rk[ds], rv[ds], rg[ds], ref_state,
                                 torch.tensor(slots[2:])).float()
    ref = torch.cat([o_meta, o_pf, o_dec]).reshape(T, H * D)

    # ---- module on the kernels
    core = HLinearAttention(hidden_size=Htot * D, num_heads=Htot, num_key_value_heads=Hkvtot,
                            num_meta_tokens=M, layer_idx=layer_idx, tp_size=tp, tp_rank=rank)
    d_state = state.to(DEV)
    meta_lin = torch.zeros(Hs, D, D, dtype=torch.bfloat16, device=DEV)
    params = HymetaCacheParams(d_state, meta_lin, torch.empty(0, device=DEV),
                               torch.tensor(slots, dtype=torch.int32, device=DEV))
    md = _md(2, lens, n_dec, ctx + [0] * n_dec)
    out = core.forward(qkv.to(DEV), params, md, has_meta_cache=False,
                       lower_bound=lb_full.to(DEV) if use_lb else None)
    torch.cuda.synchronize()
    assert out.shape == (T, H * D)
    assert_close(out, ref, "HLinearAttention mixed batch")
    assert_close(meta_lin, S_meta, "meta_linear_cache")
    assert_close(d_state, ref_state, "slot states")
    for s in (1, 4, 6):
        assert torch.equal(d_state[s].cpu(), state[s])

    # ---- steady state: meta cache present, decode only
    qkv2 = _qkv(n_dec, H, Hkv, 13)
    q2 = qkv2[:, :H * D].view(n_dec, H, D)
    k2 = qkv2[:, H * D:(H + Hkv) * D].view(n_dec, Hkv, D)
    v2 = qkv2[:, (H + Hkv) * D:].view(n_dec, Hkv, D)
    if use_lb:
        rq, rk, rg = ogla.gate_70b(q2, k2.repeat_interleave(H // Hkv, 1), v2, lower_bound=lb_full,
                                   layer_idx=layer_idx, num_heads=Htot, head_start=rank * H)
        rv = v2.repeat_interleave(H // Hkv, 1)
    else:
        rq, rk, rg = ogla.gate_70b(q2, k2, v2)
        rv = v2
    ref2 = ogla.gla_decode_step(rq, rk, rv, rg, ref_state, torch.tensor(slots[2:]))
    params.state_indices_tensor = torch.tensor(slots[2:], dtype=torch.int32, device=DEV)
    out2 = core.forward(qkv2.to(DEV), params, _md(0, [], n_dec, [0] * n_dec), has_meta_cache=True,
                        lower_bound=lb_full.to(DEV) if use_lb else None)
    torch.cuda.synchronize()
    assert_close(out2, ref2.reshape(n_dec, H * D), "HLinearAttention decode only")
    assert_close(d_state, ref_state, "slot states after decode")


def test_gated_linear_attention_7b_mixed_batch(ops):
    from vllm_hymeta_b200.model_for_7B.gla_attention import GatedLinearAttention
    from vllm_hymeta_b200.model_for_7B.gla_cache import GLACacheParams
    H, Hkv = 28, 4
    lens, n_dec = [45, 130], 2
    T = sum(lens) + n_dec
    qkv = _qkv(T, H, Hkv, 21)
    gen = torch.Generator().manual_seed(22)
    gk = (torch.randn(T, Hkv * D, generator=gen) * 2).bfloat16()
    w = (1 + 0.1 * torch.randn(D, generator=gen)).bfloat16()
    state = (torch.randn(6, Hkv, D, D, generator=gen) * 0.2).bfloat16()
    slots = [4, 1, 0, 5]
    q = qkv[:, :H * D].view(T, H, D)
    k = qkv[:, H * D:(H + Hkv) * D].view(T, Hkv, D)
    v = qkv[:, (H + Hkv) * D:].view(T, Hkv, D)
    rq, rk, rg = ogla.gate_7b(q, k, gk.view(T, Hkv, D))
    ref_state = state.clone()
    n_pf = sum(lens)
    # request 0 is NEW (context 0) on a recycled, non-zero slot: it must start from zero
    # (modeling_gla_swa.py:391-414); request 1 continues a chunked prefill from its slot state
    ctx = [0, 64, 200, 300]
    o_pf = ogla.gla_prefill_varlen(rq[:n_pf], rk[:n_pf], v[:n_pf], rg[:n_pf], ref_state,
                                   [0, lens[0], n_pf], slots[:2], [ogla.INIT_ZERO, ogla.INIT_SLOT])
    o_dec = ogla.gla_decode_step(rq[n_pf:], rk[n_pf:], v[n_pf:], rg[n_pf:], ref_state,
                                 torch.tensor(slots[2:]))
    ref = ogla.rms_norm_head(torch.cat([o_pf, o_dec]), w, 1e-6).reshape(T, H * D)

    core = GatedLinearAttention(H * D, H, Hkv, norm_eps=1e-6, g_norm_weight=w.to(DEV))
    d_state = state.to(DEV)
    params = GLACacheParams(d_state, torch.tensor(slots, dtype=torch.int32, device=DEV))
    out = core.forward(qkv.to(DEV), gk.to(DEV), params, _md(2, lens, n_dec, ctx))
    torch.cuda.synchronize()
    # the norm divides by the row rms, so compare with a slightly wider max-abs band
    assert_close(out, ref, "GatedLinearAttention mixed batch", maxabs_frac=2 ** -6, rel_rms=8e-3)
    assert_close(d_state, ref_state, "7B slot states")


def test_gate_7b_folded_into_gla_kernels(ops):
    """f2: raw q / k / gk rows through the fused paths give BIT-IDENTICAL outputs and states to
    ``gla_gate_7b`` followed by the plain kernels (relu(x) * e == max(x * e, 0); the decode kernel
    repeats the gate kernel's logsigmoid arithmetic)."""
    H, Hkv = 28, 4
    lens = [70, 33, 200]
    T, B = sum(lens), 37
    qkv = _qkv(T + B, H, Hkv, 31).to(DEV)
    gen = torch.Generator().manual_seed(32)
    gk = (torch.randn(T + B, Hkv, D, generator=gen) * 3).bfloat16().to(DEV)
    w = (1 + 0.1 * torch.randn(D, generator=gen)).bfloat16().to(DEV)
    state = (torch.randn(B + 4, Hkv, D, D, generator=gen) * 0.2).bfloat16().to(DEV)
    q = qkv[:, :H * D].view(-1, H, D)
    k = qkv[:, H * D:(H + Hkv) * D].view(-1, Hkv, D)
    v = qkv[:, (H + Hkv) * D:].view(-1, Hkv, D)
    rq, rk, rg = ops.gla_gate_7b(q, k, gk)
    assert torch.equal(ops.gla_gate_7b_g(gk), rg)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    cu = i32([0, lens[0], lens[0] + lens[1], T])
    sl, md = i32([1, -1, 3]), i32([1, 0, 1])
    for gw in (None, w):
        s1, s2 = state.clone(), state.clone()
        o1 = ops.gla_prefill_varlen(rq[:T], rk[:T], v[:T], rg[:T], s1, cu, sl, md, gnorm_weight=gw)
        o2 = ops.gla_prefill_varlen(q[:T], k[:T], v[:T], rg[:T], s2, cu, sl, md, gnorm_weight=gw, relu_qk=True)
        assert torch.equal(o1, o2) and torch.equal(s1, s2)
        # the gate looked up by the kernel's gate warps from the table of all 65536 bf16 inputs
        s3 = state.clone()
        o3 = ops.gla_prefill_varlen(q[:T], k[:T], v[:T], gk[:T], s3, cu, sl, md, gnorm_weight=gw, relu_qk=True,
                                    gate_table=ops.gla_gate_7b_table(16.0))
        assert torch.equal(o1, o3) and torch.equal(s1, s3)
        slots = torch.arange(B, dtype=torch.int32, device=DEV) + 2
        slots[5] = -1
        d1 = ops.gla_decode(rq[T:], rk[T:], v[T:], rg[T:], s1, slots, gnorm_weight=gw)
        d2 = ops.gla_decode(q[T:], k[T:], v[T:], gk[T:], s2, slots, gnorm_weight=gw, gate_normalizer=16.0)
        assert torch.equal(d1, d2) and torch.equal(s1, s2)


def test_prefill_epilogues_emit_sumsq_partials(ops):
    """f1: both prefill kernels can emit, per (output row, head), the sum of squares of the bf16 row
    they store -- the branch merge's RMS-norm partials.  Checked against the same sum taken in torch
    from the kernels' own outputs (fp32 summation order differs: relative 1e-5)."""
    H, Hkv, M = 14, 2, 128
    lens = [300, 70, 513]
    T = sum(lens)
    gen = torch.Generator().manual_seed(41)
    cu = torch.tensor([0, 300, 370, T], dtype=torch.int32, device=DEV)
    q = torch.randn(T, H, D, generator=gen).bfloat16().to(DEV)
    k = torch.randn(T, Hkv, D, generator=gen).bfloat16().to(DEV)
    v = torch.randn(T, Hkv, D, generator=gen).bfloat16().to(DEV)
    mk = torch.randn(M, Hkv, D, generator=gen).bfloat16().to(DEV)
    mv = torch.randn(M, Hkv, D, generator=gen).bfloat16().to(DEV)
    q2 = torch.randn(M, H, D, generator=gen).bfloat16().to(DEV)
    ss = torch.full((M + T, H), -1.0, dtype=torch.float32, device=DEV)
    o = ops.swa_meta_prefill_varlen(q, k, v, q2, mk, mv, cu, cu, max(lens), 4095, sumsq_out=ss)
    ref = o.float().pow(2).sum(-1)
    assert torch.allclose(ss, ref, rtol=1e-5, atol=1e-6), (ss - ref).abs().max().item()
    # GLA prefill (70B-style fp32 gates, meta-state init)
    qg, kg, vg, gg = ogla.synth_gla_inputs(T, H, Hkv, D, "70b", seed=42)
    state = torch.zeros(4, Hkv, D, D, dtype=torch.bfloat16, device=DEV)
    ss2 = torch.full((T, H), -1.0, dtype=torch.float32, device=DEV)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    og = ops.gla_prefill_varlen(qg.to(DEV), kg.to(DEV), vg.to(DEV), gg.to(DEV), state, cu, i32([0, 1, 2]),
                                i32([0, 0, 0]), sumsq_out=ss2)
    ref2 = og.float().pow(2).sum(-1)
    assert torch.allclose(ss2, ref2, rtol=1e-5, atol=1e-6), (ss2 - ref2).abs().max().item()


# ------------------------------------------------------------------------------------- a7-a10 (backend)
def test_flash_attention_backend_warmup_then_decode(ops):
    from vllm_hymeta_b200.attention.meta_attention import MetaAttentionMetadataBuilder, SeqData
    from vllm_hymeta_b200.models.attention import FlashAttention
    from vllm_hymeta_b200.models.hymeta_cache import HymetaCacheParams
    H, Hkv, M, bs, W = 14, 2, 128, 16, 64
    lens = [150, 40]
    nb = 40
    gen = torch.Generator().manual_seed(31)
    T = M + sum(lens)
    q = torch.randn(T, H * D, generator=gen).bfloat16()
    k = torch.randn(T, Hkv * D, generator=gen).bfloat16()
    v = torch.randn(T, Hkv * D, generator=gen).bfloat16()
    perm = torch.randperm(nb, generator=gen).tolist()
    tables = [perm[:11], perm[11:15]]             # room for decode growth
    kv_cache = torch.randn(2, nb, bs, Hkv, D, generator=gen).bfloat16()      # garbage everywhere
    layer = FlashAttention(H, Hkv, D, num_meta_tokens=M, sliding_window=W)
    d_cache = kv_cache.to(DEV)
    layer.attn.bind_kv_cache(d_cache)
    meta_fa = torch.zeros(2, M, Hkv, D, dtype=torch.bfloat16, device=DEV)
    params = HymetaCacheParams(torch.empty(0), torch.empty(0), meta_fa, torch.empty(0))

    # ---- pass 1: [meta | two prompts], nothing cached yet
    b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device=DEV)
    for L, bt in zip(lens, tables):
        b.add(SeqData(True, L, L, 0, bt))
    md = b.build()
    out = layer.forward(q.to(DEV), k.to(DEV), v.to(DEV), params, md, has_meta_cache=False)
    torch.cuda.synchronize()
    q3, k3, v3 = q.view(T, H, D), k.view(T, Hkv, D), v.view(T, Hkv, D)
    ref = oswa.swa_meta_prefill_varlen(q3[M:], k3[M:], v3[M:], q3[:M], k3[:M], v3[:M],
                                       [0, lens[0], sum(lens)], [0, lens[0], sum(lens)], W - 1,
                                       out_dtype=torch.float32)
    assert_close(out.view(T, H, D), ref, "FlashAttention warm-up pass (meta rows first)")
    assert torch.equal(meta_fa[0].cpu(), k3[:M]) and torch.equal(meta_fa[1].cpu(), v3[:M])
    ref_cache = kv_cache.clone()
    oswa.kv_write_paged(k3[M:], v3[M:], ref_cache[0], ref_cache[1], md.slot_mapping.cpu().tolist())
    assert torch.equal(d_cache.cpu(), ref_cache), "paged KV write must be bit-exact"

    # ---- pass 2: one decode token per request, meta caches warm
    qd = torch.randn(2, H * D, generator=gen).bfloat16()
    kd = torch.randn(2, Hkv * D, generator=gen).bfloat16()
    vd = torch.randn(2, Hkv * D, generator=gen).bfloat16()
    b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device=DEV)
    for L, bt in zip(lens, tables):
        b.add(SeqData(False, L + 1, 1, L, bt))
    md2 = b.build()
    out2 = layer.forward(qd.to(DEV), kd.to(DEV), vd.to(DEV), params, md2, has_meta_cache=True)
    torch.cuda.synchronize()
    oswa.kv_write_paged(kd.view(2, Hkv, D), vd.view(2, Hkv, D), ref_cache[0], ref_cache[1],
                        md2.slot_mapping.cpu().tolist())
    assert torch.equal(d_cache.cpu(), ref_cache)
    bt_rows = md2.block_tables.cpu().tolist()
    ref2 = oswa.swa_meta_decode_paged(qd.view(2, H, D), ref_cache[0], ref_cache[1], bt_rows,
                                      [L + 1 for L in lens], k3[:M], v3[:M], W - 1,
                                      out_dtype=torch.float32)
    assert_close(out2.view(2, H, D), ref2, "FlashAttention decode pass")


def test_flash_attention_backend_prefix_hit_chunked_and_multitoken_decode(ops):
    """``MetaAttentionImpl.forward`` branches that read keys from the paged cache: prefix-cache hit /
    chunked prefill mixed with a decode row (reference meta_attention.py:886-919), then multi-token
    decode (``max_decode_query_len > 1``, :924-953).  Pages are written by the same forward."""
    from vllm_hymeta_b200.attention.meta_attention import MetaAttentionMetadataBuilder, SeqData
    from vllm_hymeta_b200.models.attention import FlashAttention
    from vllm_hymeta_b200.models.hymeta_cache import HymetaCacheParams
    H, Hkv, M, bs, W = 14, 2, 128, 16, 64
    ctx = [150, 40]                       # already cached (filled by a first no-cache prefill)
    nb = 48
    gen = torch.Generator().manual_seed(41)
    perm = torch.randperm(nb, generator=gen).tolist()
    tables = [perm[:16], perm[16:21], perm[21:25]]
    kv_cache = torch.full((2, nb, bs, Hkv, D), float("nan"), dtype=torch.bfloat16)   # unowned rows: NaN
    layer = FlashAttention(H, Hkv, D, num_meta_tokens=M, sliding_window=W)
    d_cache = kv_cache.to(DEV)
    layer.attn.bind_kv_cache(d_cache)
    k2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
    v2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
    meta_fa = torch.stack([k2, v2]).to(DEV)
    params = HymetaCacheParams(torch.empty(0), torch.empty(0), meta_fa, torch.empty(0))
    rnd = lambda T, h: torch.randn(T, h * D, generator=gen).bfloat16()
    ref_cache = kv_cache.clone()

    def step(seqs):
        """seqs: (is_prompt, context, new tokens, table); returns (out, ref rows builder inputs)."""
        b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device=DEV,
                                         chunked_prefill_enabled=True)
        for is_prompt, c, n, bt in seqs:
            b.add(SeqData(is_prompt, c + n, n, c, bt), prefix_cache_hit=is_prompt and c > 0)
        md = b.build()
        T = sum(n for _, _, n, _ in seqs)
        q, k, v = rnd(T, H), rnd(T, Hkv), rnd(T, Hkv)
        out = layer.forward(q.to(DEV), k.to(DEV), v.to(DEV), params, md, has_meta_cache=True)
        torch.cuda.synchronize()
        oswa.kv_write_paged(k.view(T, Hkv, D), v.view(T, Hkv, D), ref_cache[0], ref_cache[1],
                            md.slot_mapping.cpu().tolist())
        assert torch.equal(torch.nan_to_num(d_cache.cpu().float(), nan=7.0),
                           torch.nan_to_num(ref_cache.float(), nan=7.0)), "paged KV write must be bit-exact"
        return out.view(T, H, D), q.view(T, H, D), md

    # ---- pass 1: plain prompts (no cache hit) fill the pages
    step([(True, 0, ctx[0], tables[0]), (True, 0, ctx[1], tables[1])])

    # ---- pass 2: [chunk of request 0 on top of its 150 cached tokens | a new prompt | decode of request 1]
    new0, new2 = 70, 33
    out, q, md = step([(True, ctx[0], new0, tables[0]), (True, 0, new2, tables[2]),
                       (False, ctx[1], 1, tables[1])])
    assert md.num_prefills == 2 and md.num_decode_tokens == 1
    bt_rows = md.block_tables.cpu().tolist()
    ref_p = oswa.swa_meta_prefill_paged(q[:new0 + new2], ref_cache[0], ref_cache[1], bt_rows[:2],
                                        [0, new0, new0 + new2], [ctx[0] + new0, new2], k2, v2, W - 1,
                                        out_dtype=torch.float32)
    ref_d = oswa.swa_meta_decode_paged(q[new0 + new2:], ref_cache[0], ref_cache[1], bt_rows[2:],
                                       [ctx[1] + 1], k2, v2, W - 1, out_dtype=torch.float32)
    assert_close(out, torch.cat([ref_p, ref_d]), "prefix-hit + chunked prefill + decode")

    # ---- pass 3: multi-token decode (3 and 2 new tokens on requests 0 and 1)
    L0, L1 = ctx[0] + new0, ctx[1] + 1
    out, q, md = step([(False, L0, 3, tables[0]), (False, L1, 2, tables[1])])
    assert md.max_decode_query_len == 3 and md.num_prefills == 0
    bt_rows = md.block_tables.cpu().tolist()
    ref = oswa.swa_meta_prefill_paged(q, ref_cache[0], ref_cache[1], bt_rows, [0, 3, 5],
                                      [L0 + 3, L1 + 2], k2, v2, W - 1, out_dtype=torch.float32)
    assert_close(out, ref, "multi-token decode over the paged cache")


# ------------------------------------------------------------------------------------- f3 (fused prologue)
def test_flash_attention_fused_prologue_equals_unfused(ops):
    """``forward_fused(qkv, positions)`` == ``rotary_emb(new_positions, q, k)`` + ``forward``
    (modeling_hymeta.py:757-796 with the :1314-1324 position shift): output, pages and meta
    caches bit-identical, on the warm-up pass and on a decode pass."""
    from vllm_hymeta_b200.attention.meta_attention import MetaAttentionMetadataBuilder, SeqData
    from vllm_hymeta_b200.models.attention import FlashAttention, HymetaRotaryEmbedding
    from vllm_hymeta_b200.models.hymeta_cache import HymetaCacheParams
    H, Hkv, M, bs, W = 14, 2, 128, 16, 64
    lens = [150, 40]
    nb = 40
    gen = torch.Generator().manual_seed(77)
    perm = torch.randperm(nb, generator=gen).tolist()
    tables = [perm[:11], perm[11:15]]
    kv_cache = torch.randn(2, nb, bs, Hkv, D, generator=gen).bfloat16()
    rot = HymetaRotaryEmbedding(D, D, 4096, 10000, True, torch.bfloat16)

    def run(fused):
        layer = FlashAttention(H, Hkv, D, num_meta_tokens=M, sliding_window=W)
        cache = kv_cache.to(DEV)
        layer.attn.bind_kv_cache(cache)
        meta_fa = torch.zeros(2, M, Hkv, D, dtype=torch.bfloat16, device=DEV)
        params = HymetaCacheParams(torch.empty(0), torch.empty(0), meta_fa, torch.empty(0))
        g2 = torch.Generator().manual_seed(78)
        outs = []
        for decode in (False, True):
            b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device=DEV)
            for L, bt in zip(lens, tables):
                b.add(SeqData(False, L + 1, 1, L, bt) if decode else SeqData(True, L, L, 0, bt))
            md = b.build()
            T = 2 if decode else M + sum(lens)
            qkv = torch.randn(T, (H + 2 * Hkv) * D, generator=g2).bfloat16().to(DEV)
            pos = (torch.tensor(lens) if decode else
                   torch.cat([torch.arange(L) for L in lens])).to(DEV)
            if fused:
                out = layer.forward_fused(qkv, pos, params, md, rot, has_meta_cache=decode)
            else:
                q, k, v = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)
                new_pos = pos + M
                if not decode:
                    new_pos = torch.cat([torch.arange(M, device=DEV), new_pos])
                q, k = rot(new_pos, q, k)
                out = layer.forward(q, k, v, params, md, has_meta_cache=decode)
            torch.cuda.synchronize()
            outs.append(out.cpu())
        return outs, cache.cpu(), meta_fa.cpu()

    (a1, a2), ca, ma = run(False)
    (b1, b2), cb, mb = run(True)
    assert torch.isfinite(a1.float()).all() and torch.isfinite(a2.float()).all()
    assert torch.equal(a1, b1) and torch.equal(a2, b2)
    assert torch.equal(ca, cb) and torch.equal(ma, mb)
    # and the rotation itself is the oracle's, bit for bit
    x = torch.randn(33, H * D, generator=gen).bfloat16()
    kx = torch.randn(33, Hkv * D, generator=gen).bfloat16()
    p = torch.randint(0, 4096, (33,), generator=gen)
    rq, rk = rot(p.to(DEV), x.to(DEV), kx.to(DEV))
    tab = oswa.rope_cos_sin_cache(D, 4096, 10000)
    assert torch.equal(rq.cpu().view(33, H, D), oswa.rope_neox_table(x.view(33, H, D), p, tab))
    assert torch.equal(rk.cpu().view(33, Hkv, D), oswa.rope_neox_table(kx.view(33, Hkv, D), p, tab))


# ------------------------------------------------------------------------------------- f2 (g_norm epilogue)
def _assert_ulp_close(got, ref, what, max_frac=0.02, floor=0.0):
    """Same bf16 arithmetic up to the fp32 summation order of the variance and rsqrt rounding: a
    flipped first rounding (x * rstd) is one bf16 ulp, which the multiplication by w can turn into
    two ulps of the result -- every element within that, all but a small fraction bit-identical."""
    got, ref = got.float().cpu(), ref.float().cpu()
    assert torch.isfinite(got).all(), what
    diff = (got - ref).abs()
    # floor: magnitude of the addends when the result is a sum that may cancel (branch merge)
    bad = diff > 2.0 ** -6 * ref.abs().clamp(min=floor) + 1e-30
    assert not bad.any(), (f"{what}: more than two bf16 ulps apart at {int(bad.sum())} elements, "
                           f"worst {diff.max().item():.4g}")
    frac = (diff > 0).float().mean().item()
    assert frac <= max_frac, f"{what}: {frac:.3%} of elements differ"


@pytest.mark.parametrize("H,Hkv,gdt", [(28, 4, torch.bfloat16), (7, 1, torch.bfloat16),
                                        (8, 8, torch.float32), (6, 2, torch.bfloat16)])
def test_gnorm_epilogue_decode(ops, H, Hkv, gdt):
    """Fused g_norm == the kernel's own bf16 output followed by the reference RMSNorm arithmetic
    (gla_attention.py:233), and the state update is bit-identical with and without the epilogue."""
    B = 37
    gen = torch.Generator().manual_seed(H)
    q = torch.randn(B, H, D, generator=gen).relu().bfloat16().to(DEV)
    k = torch.randn(B, Hkv, D, generator=gen).relu().bfloat16().to(DEV)
    v = torch.randn(B, Hkv, D, generator=gen).bfloat16().to(DEV)
    g = (-torch.rand(B, Hkv, D, generator=gen) * 0.1).to(gdt).to(DEV)
    w = (1 + 0.1 * torch.randn(D, generator=gen)).bfloat16()
    state = (torch.randn(B + 2, Hkv, D, D, generator=gen) * 0.2).bfloat16()
    slots = torch.randperm(B + 2, generator=gen)[:B].to(torch.int32)
    slots[3] = -1
    s_a, s_b = state.to(DEV), state.to(DEV)
    plain = ops.gla_decode(q, k, v, g, s_a, slots.to(DEV))
    fused = ops.gla_decode(q, k, v, g, s_b, slots.to(DEV), gnorm_weight=w.to(DEV), gnorm_eps=1e-6)
    torch.cuda.synchronize()
    assert torch.equal(s_a, s_b)
    _assert_ulp_close(fused, ogla.rms_norm_head(plain.cpu(), w, 1e-6), f"decode g_norm H={H}/{Hkv}")
    assert (fused[3] == 0).all()                                  # PAD slot row stays zero


@pytest.mark.parametrize("lens,seg_len", [([100, 333, 16, 1], 0), ([1500], 256)])
def test_gnorm_epilogue_prefill(ops, lens, seg_len):
    H, Hkv, T, N = 28, 4, sum(lens), len(lens)
    q, k, v, g = ogla.synth_gla_inputs(T, H, Hkv, D, "7b", seed=T)
    gen = torch.Generator().manual_seed(3)
    w = (1 + 0.1 * torch.randn(D, generator=gen)).bfloat16()
    cu = [0]
    for L in lens:
        cu.append(cu[-1] + L)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    args = [x.to(DEV) for x in (q, k, v, g)]
    s_a = torch.zeros(N, Hkv, D, D, dtype=torch.bfloat16, device=DEV)
    s_b = torch.zeros_like(s_a)
    kw = dict(max_seqlen=max(lens), seg_len=seg_len)
    plain = ops.gla_prefill_varlen(*args, s_a, i32(cu), i32(list(range(N))), i32([0] * N), **kw)
    fused = ops.gla_prefill_varlen(*args, s_b, i32(cu), i32(list(range(N))), i32([0] * N),
                                   gnorm_weight=w.to(DEV), gnorm_eps=1e-6, **kw)
    torch.cuda.synchronize()
    assert torch.equal(s_a, s_b)
    _assert_ulp_close(fused, ogla.rms_norm_head(plain.cpu(), w, 1e-6), f"prefill g_norm {lens}")


# ------------------------------------------------------------------------------------- f1 (branch merge)
@pytest.mark.parametrize("T,hidden,tp", [(777, 3584, 1), (64, 3584, 1), (300, 3584, 2), (129, 3584, 4),
                                         (1, 128, 1)])
def test_branch_norm_merge(ops, T, hidden, tp):
    """norm1(out_attn), norm2(out_linear), average, bf16 cast (modeling_hymeta.py:909-919) in one
    pass; tp > 1 simulates the head shards on one GPU with the cross-shard sum of the per-token
    means done here (the all-reduce of the real run)."""
    gen = torch.Generator().manual_seed(T + tp)
    xa = (torch.randn(T, hidden, generator=gen) * 3).bfloat16()
    xl = (torch.randn(T, hidden, generator=gen) * 0.2).bfloat16()
    w1 = 1 + 0.1 * torch.randn(hidden, generator=gen)
    w2 = 1 + 0.1 * torch.randn(hidden, generator=gen)
    eps, hl = 1e-6, hidden // tp
    sl = [slice(r * hl, (r + 1) * hl) for r in range(tp)]
    tot_a = sum(xa[:, s].float().pow(2).mean(-1, keepdim=True, dtype=torch.float32) for s in sl)
    tot_l = sum(xl[:, s].float().pow(2).mean(-1, keepdim=True, dtype=torch.float32) for s in sl)
    d_means = [ops.branch_sumsq(xa[:, s].to(DEV), xl[:, s].to(DEV)) for s in sl]
    red = sum(d_means)                                     # what dist.all_reduce(SUM) leaves on every rank
    for r, s in enumerate(sl):
        ref = ogla.branch_norm_merge(xa[:, s], xl[:, s], w1[s], w2[s], eps, tp,
                                     (lambda v: tot_a, lambda v: tot_l))
        got = ops.branch_norm_merge(xa[:, s].to(DEV), xl[:, s].to(DEV), w1[s].to(DEV).contiguous(),
                                    w2[s].to(DEV).contiguous(), eps, tp_world=tp,
                                    means=red if tp > 1 else None)
        torch.cuda.synchronize()
        # the two normalised addends are O(1) and may cancel: ulps are those of the addends
        _assert_ulp_close(got, ref, f"branch merge T={T} tp={tp} rank={r}", floor=1.0)
    # in place over one input
    x_in = xa[:, sl[0]].to(DEV).contiguous()
    keep = x_in.clone()
    ops.branch_norm_merge(x_in, xl[:, sl[0]].to(DEV), w1[sl[0]].to(DEV).contiguous(),
                          w2[sl[0]].to(DEV).contiguous(), eps, tp_world=tp,
                          means=red if tp > 1 else None, out=x_in)
    got2 = ops.branch_norm_merge(keep, xl[:, sl[0]].to(DEV), w1[sl[0]].to(DEV).contiguous(),
                                 w2[sl[0]].to(DEV).contiguous(), eps, tp_world=tp,
                                 means=red if tp > 1 else None)
    torch.cuda.synchronize()
    assert torch.equal(x_in, got2)
    with pytest.raises(ValueError):
        ops.branch_norm_merge(keep, keep, w1[:5].to(DEV), w2[sl[0]].to(DEV).contiguous(), eps)


def test_branch_norm_merge_edge_cases(ops):
    e = torch.empty(0, 3584, dtype=torch.bfloat16, device=DEV)
    w = torch.ones(3584, device=DEV)
    assert ops.branch_norm_merge(e, e, w, w, 1e-6).shape == (0, 3584)          # empty batch: no-op
    assert ops.branch_sumsq(e, e).shape == (2, 0)
    x = torch.randn(4, 100, device=DEV).bfloat16()                              # hidden not a multiple of 8
    with pytest.raises(ValueError):
        ops.branch_norm_merge(x, x, torch.ones(100, device=DEV), torch.ones(100, device=DEV), 1e-6)
    with pytest.raises(ValueError):
        ops.branch_norm_merge(e.cpu(), e.cpu(), w, w, 1e-6)                     # no CPU path
    # wide rows (> 3584) take the streaming kernel: same arithmetic
    gen = torch.Generator().manual_seed(8)
    xa = torch.randn(33, 8192, generator=gen).bfloat16()
    xl = torch.randn(33, 8192, generator=gen).bfloat16()
    w1, w2 = 1 + 0.1 * torch.randn(8192, generator=gen), 1 + 0.1 * torch.randn(8192, generator=gen)
    ref = ogla.branch_norm_merge(xa, xl, w1, w2, 1e-5)
    got = ops.branch_norm_merge(xa.to(DEV), xl.to(DEV), w1.to(DEV), w2.to(DEV), 1e-5)
    torch.cuda.synchronize()
    _assert_ulp_close(got, ref, "branch merge, wide rows", floor=1.0)


# ------------------------------------------------------------------------------------- b (boundary)
def test_platform_binds_to_vllm_nvml_platform():
    """On a GPU host the platform class derives from the installed vLLM's NVML CUDA platform
    (reference platform.py:7-10) and answers the V1 selection protocol."""
    pytest.importorskip("vllm")
    from types import SimpleNamespace
    from vllm_hymeta_b200.platform import HymetaCudaPlatform as P
    assert P.__mro__[1].__name__ == "NvmlCudaPlatform"
    cfg = SimpleNamespace(use_mla=False, head_size=128)
    assert P.get_attn_backend_cls(None, cfg, 28).endswith("meta_attention_v1.MetaAttentionV1Backend")


def test_custom_op_path_equals_direct_call(ops):
    """``torch.ops.vllm.unified_meta_attention_with_output`` (reference models/attention.py:361-387)
    through the forward context == the direct ``impl.forward`` call, bit for bit."""
    from vllm_hymeta_b200.attention.meta_attention import MetaAttentionMetadataBuilder, SeqData
    from vllm_hymeta_b200.models import attention as A
    H, Hkv, M, bs, W = 14, 2, 128, 16, 64
    lens = [90, 33]
    gen = torch.Generator().manual_seed(51)
    nb = 20
    perm = torch.randperm(nb, generator=gen).tolist()
    tables = [perm[:7], perm[7:11]]
    T = sum(lens)
    q = torch.randn(T, H * D, generator=gen).bfloat16().to(DEV)
    k = torch.randn(T, Hkv * D, generator=gen).bfloat16().to(DEV)
    v = torch.randn(T, Hkv * D, generator=gen).bfloat16().to(DEV)
    k2 = torch.randn(M, Hkv, D, generator=gen).bfloat16().to(DEV)
    v2 = torch.randn(M, Hkv, D, generator=gen).bfloat16().to(DEV)
    outs = []
    for direct in (True, False):
        name = f"parity.layers.{int(direct)}.attn"
        layer = A.MetaAttention(H, D, D ** -0.5, Hkv, per_layer_sliding_window=W, num_meta_tokens=M,
                                prefix=name, use_direct_call=direct)
        layer.bind_kv_cache(torch.zeros(2, nb, bs, Hkv, D, dtype=torch.bfloat16, device=DEV))
        b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device=DEV)
        for L, bt in zip(lens, tables):
            b.add(SeqData(True, L, L, 0, bt))
        md = b.build()
        with A.set_forward_context(md):
            outs.append(layer.forward(q, k, v, key2=k2, value2=v2).clone())
        del A._LAYERS[name]
    torch.cuda.synchronize()
    assert torch.isfinite(outs[0].float()).all() and torch.equal(outs[0], outs[1])


# ------------------------------------------------------------------------------------- f4 (V1 adapter)
def test_v1_adapter_mixed_batch_vs_oracle(ops):
    """``MetaAttentionV1Impl`` driven the way the installed vLLM's V1 runner drives a backend:
    ``do_kv_cache_update`` then ``forward`` on a reordered [decodes | prefills / extends] batch with
    a padded query tensor, against the oracle over the paged cache; with and without meta K/V."""
    from vllm_hymeta_b200.attention.meta_attention_v1 import (MetaAttentionV1Backend,
                                                              MetaAttentionV1Impl,
                                                              MetaAttentionV1MetadataBuilder)
    H, Hkv, bs, W, M = 14, 2, 16, 64, 128
    # request: (context, new tokens); two decodes first, then an extend (prefix hit) and a new prompt
    reqs = [(40, 1), (300, 1), (150, 70), (0, 33)]
    gen = torch.Generator().manual_seed(61)
    nblk = [(c + n + bs - 1) // bs for c, n in reqs]
    nb = sum(nblk) + 3
    perm = torch.randperm(nb, generator=gen).tolist()
    kv = torch.full(MetaAttentionV1Backend.get_kv_cache_shape(nb, bs, Hkv, D), float("nan"),
                    dtype=torch.bfloat16)
    bt = torch.zeros(len(reqs), max(nblk), dtype=torch.int32)
    at = 0
    for i, (c, n) in enumerate(reqs):
        row = perm[at:at + nblk[i]]
        at += nblk[i]
        bt[i, :nblk[i]] = torch.tensor(row, dtype=torch.int32)
        for p in range(c):                      # the cached context
            kv[0, row[p // bs], p % bs] = torch.randn(Hkv, D, generator=gen).bfloat16()
            kv[1, row[p // bs], p % bs] = torch.randn(Hkv, D, generator=gen).bfloat16()
    T = sum(n for _, n in reqs)
    qsl = torch.tensor([0] + torch.tensor([n for _, n in reqs]).cumsum(0).tolist(), dtype=torch.int32)
    seq = torch.tensor([c + n for c, n in reqs], dtype=torch.int32)
    slots = torch.tensor([int(bt[i, p // bs]) * bs + p % bs for i, (c, n) in enumerate(reqs)
                          for p in range(c, c + n)], dtype=torch.int64)
    pad = 3                                     # CUDA-graph style padding of the token tensors
    q = torch.randn(T + pad, H * D, generator=gen).bfloat16()
    k = torch.randn(T + pad, Hkv * D, generator=gen).bfloat16()
    v = torch.randn(T + pad, Hkv * D, generator=gen).bfloat16()
    k2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
    v2 = torch.randn(M, Hkv, D, generator=gen).bfloat16()
    cm = SimpleNamespace(query_start_loc=qsl.to(DEV), query_start_loc_cpu=qsl, seq_lens=seq.to(DEV),
                         num_reqs=len(reqs), num_actual_tokens=T, max_query_len=70, max_seq_len=301,
                         block_table_tensor=bt.to(DEV), slot_mapping=slots.to(DEV), causal=True)
    md = MetaAttentionV1MetadataBuilder(SimpleNamespace(block_size=bs), ["l"], None, DEV).build(0, cm)
    ref_kv = kv.clone()
    oswa.kv_write_paged(k[:T].view(T, Hkv, D), v[:T].view(T, Hkv, D), ref_kv[0], ref_kv[1], slots.tolist())
    impl = MetaAttentionV1Impl(H, D, D ** -0.5, Hkv, None, W, "auto")
    for with_meta in (True, False):
        layer = SimpleNamespace(meta_key=k2.to(DEV) if with_meta else None,
                                meta_value=v2.to(DEV) if with_meta else None)
        d_kv = kv.to(DEV)
        out = torch.full((T + pad, H * D), 9.0, dtype=torch.bfloat16, device=DEV)
        impl.do_kv_cache_update(layer, k.to(DEV), v.to(DEV), d_kv, md.slot_mapping)
        impl.forward(layer, q.to(DEV).view(-1, H, D), k.to(DEV), v.to(DEV), d_kv, md, output=out.view(-1, H, D))
        torch.cuda.synchronize()
        assert torch.equal(torch.nan_to_num(d_kv.cpu().float(), nan=7.0),
                           torch.nan_to_num(ref_kv.float(), nan=7.0))
        mk, mv = (k2, v2) if with_meta else (None, None)
        q3 = q.view(-1, H, D)
        ref_d = oswa.swa_meta_decode_paged(q3[:2], ref_kv[0], ref_kv[1], bt[:2].tolist(), seq[:2].tolist(),
                                           mk, mv, W - 1, out_dtype=torch.float32)
        ref_p = oswa.swa_meta_prefill_paged(q3[2:T], ref_kv[0], ref_kv[1], bt[2:].tolist(), [0, 70, 103],
                                            seq[2:].tolist(), mk, mv, W - 1, out_dtype=torch.float32)
        assert_close(out[:T].view(T, H, D), torch.cat([ref_d, ref_p]), f"V1 adapter (meta={with_meta})")
        assert (out[T:] == 9.0).all(), "padded rows must not be written"
    assert impl.forward(layer, q.to(DEV), None, None, d_kv, None, output=out).abs().max().item() == 0
