"""-m gpu cross-checks against the LIVE third-party kernels the reference calls (last
row): flash_attn (reference pins 2.7.3, fa_kernel_metatoken.py:6-8) and flash-linear-attention
(reference pins 0.1, modeling_hymeta.py:59) as installed in the image (2.8.3 / 0.5.1).  They are
CHECKERS here, never the product path: each test runs our C-ABI kernel and the library on the same
bf16 inputs and compares with the tolerances of gpu_util (both sides are bf16 approximations of the
same fp32 function).  The reference's composition around flash_attn -- window pass + meta pass +
``_update_out_and_lse`` merge (fa_kernel_metatoken.py:198-285, :26-106) -- is rebuilt from the live
flash_attn calls and the oracle's merge.  A library that cannot run on this GPU skips its tests.
"""
import pytest
import torch

from oracle import gla as ogla
from oracle import swa as oswa
from gpu_util import DEV, assert_close, make_paged

pytestmark = pytest.mark.gpu
D = 128


@pytest.fixture(scope="module")
def ops():
    from vllm_hymeta_b200 import ops as o
    return o


@pytest.fixture(scope="module")
def fa():
    try:
        import flash_attn
        q = torch.randn(16, 2, D, device=DEV, dtype=torch.bfloat16)
        cu = torch.tensor([0, 16], dtype=torch.int32, device=DEV)
        flash_attn.flash_attn_varlen_func(q, q, q, cu, cu, 16, 16, causal=True)
        torch.cuda.synchronize()
        return flash_attn
    except Exception as e:                                   # no sm_100 build, import error ...
        pytest.skip(f"flash_attn not runnable here: {type(e).__name__}: {e}")


@pytest.fixture(scope="module")
def fla_gla():
    try:
        from fla.ops.gla import fused_recurrent_gla
        x = torch.randn(1, 8, 1, D, device=DEV, dtype=torch.bfloat16)
        fused_recurrent_gla(x, x, x, gk=-x.float().abs() * 0.01)
        torch.cuda.synchronize()
        return fused_recurrent_gla
    except Exception as e:
        pytest.skip(f"fla not runnable here: {type(e).__name__}: {e}")


def _cu(lens):
    c = [0]
    for L in lens:
        c.append(c[-1] + L)
    return c


@pytest.mark.parametrize("lens,W,H,Hkv", [([300, 1000, 77], 255, 28, 4), ([4500], 4096, 7, 1),
                                          ([513, 64], -1, 8, 8)])
def test_swa_prefill_vs_flash_attn(ops, fa, lens, W, H, Hkv):
    """No meta prefix (7B): flash_attn_varlen_func(causal, window=(W,0)) is the reference call."""
    T = sum(lens)
    gen = torch.Generator().manual_seed(T)
    q, k, v = (torch.randn(T, h, D, generator=gen).bfloat16().to(DEV) for h in (H, Hkv, Hkv))
    cu = torch.tensor(_cu(lens), dtype=torch.int32, device=DEV)
    ref = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), softmax_scale=D ** -0.5,
                                    causal=True, window_size=(W, 0) if W >= 0 else (-1, -1))
    out = ops.swa_meta_prefill_varlen(q, k, v, None, None, None, cu, cu, max(lens), W)
    torch.cuda.synchronize()
    assert_close(out, ref, f"swa prefill vs flash_attn {lens} W={W}")


@pytest.mark.parametrize("lens,W,M", [([300, 1000, 77], 255, 128), ([2100], 1000, 100)])
def test_swa_meta_prefill_vs_flash_attn_two_pass(ops, fa, lens, W, M):
    """70B: the reference's two flash_attn passes + LSE merge, rebuilt on the live library, against
    our single-pass kernel (meta keys folded into the same online softmax)."""
    H, Hkv, T = 28, 4, sum(lens)
    gen = torch.Generator().manual_seed(T + M)
    q, k, v = (torch.randn(T, h, D, generator=gen).bfloat16().to(DEV) for h in (H, Hkv, Hkv))
    k2, v2 = (torch.randn(M, Hkv, D, generator=gen).bfloat16().to(DEV) for _ in range(2))
    cu = torch.tensor(_cu(lens), dtype=torch.int32, device=DEV)
    sc = D ** -0.5
    o1, lse1, _ = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), softmax_scale=sc,
                                            causal=True, window_size=(W, 0), return_attn_probs=True)
    # meta pass: every real query sees all M meta keys (one "sequence" of T queries x M keys)
    cq = torch.tensor([0, T], dtype=torch.int32, device=DEV)
    ck = torch.tensor([0, M], dtype=torch.int32, device=DEV)
    o2, lse2, _ = fa.flash_attn_varlen_func(q, k2, v2, cq, ck, T, M, softmax_scale=sc, causal=False,
                                            return_attn_probs=True)
    # lse: [H, T] (flash_attn >= 2.6 varlen layout, the index comment at fa_kernel_metatoken.py:254)
    ref, _ = oswa.merge_out_lse(o1.float().cpu()[None], lse1.float().cpu()[None],
                                o2.float().cpu()[None], lse2.float().cpu()[None])
    out = ops.swa_meta_prefill_varlen(q, k, v, None, k2, v2, cu, cu, max(lens), W)
    torch.cuda.synchronize()
    assert_close(out, ref.view(T, H, D), f"swa+meta prefill vs flash_attn two-pass {lens}")


@pytest.mark.parametrize("lens,W", [([5000, 4096, 4097, 8704, 129], 4096), ([300, 1023, 1025], -1)])
def test_swa_decode_vs_flash_attn_kvcache(ops, fa, lens, W):
    """flash_attn_with_kvcache over the 256-token block table (fa_kernel_metatoken.py:84)."""
    H, Hkv, bs, B = 28, 4, 256, len(lens)
    kc, vc, tables, _, _ = make_paged(lens, Hkv, D, bs, seed=sum(lens))
    gen = torch.Generator().manual_seed(9)
    q = torch.randn(B, H, D, generator=gen).bfloat16().to(DEV)
    kc, vc, tables = kc.to(DEV), vc.to(DEV), tables.to(DEV)
    sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
    ref = fa.flash_attn_with_kvcache(q.unsqueeze(1), kc, vc, cache_seqlens=sl, block_table=tables,
                                     softmax_scale=D ** -0.5, causal=True,
                                     window_size=(W, 0) if W >= 0 else (-1, -1))
    out = ops.swa_meta_decode_paged(q, kc, vc, tables, sl, None, None, W)
    torch.cuda.synchronize()
    assert_close(out, ref.squeeze(1), f"swa decode vs flash_attn_with_kvcache {lens} W={W}")


@pytest.mark.parametrize("model", ["7b", "70b"])
def test_gla_prefill_vs_fla_recurrent(ops, fla_gla, model):
    """fla's fused_recurrent_gla (fp32 recurrence in Triton) on the repeat_kv'ed inputs -- the
    definition the reference's fused_chunk_gla call sites rely on (modeling_hymeta.py:463,512)."""
    T, H, Hkv = 700, 28, 4
    q, k, v, g = ogla.synth_gla_inputs(T, H, Hkv, D, model, seed=21)
    G = H // Hkv
    rep = lambda x: x.repeat_interleave(G, dim=1)
    qd, kd, vd, gd = (x.to(DEV) for x in (q, k, v, g))
    ref, ref_S = fla_gla(qd[None], rep(kd)[None], rep(vd)[None], gk=rep(gd)[None].float(),
                         scale=D ** -0.5, output_final_state=True)
    state = torch.zeros(1, Hkv, D, D, dtype=torch.bfloat16, device=DEV)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    out = ops.gla_prefill_varlen(qd, kd, vd, gd, state, i32([0, T]), i32([0]), i32([0]),
                                 max_seqlen=T)
    torch.cuda.synchronize()
    assert_close(out, ref[0], f"gla prefill vs fla fused_recurrent_gla ({model})")
    assert_close(state[0], ref_S[0, ::G], f"gla final state vs fla ({model})")


# ------------------------------------------------------------------------------------------------
# The reference's OWN Triton decode kernel, run live (unmodified file from baseline/_ref, installed
# by oracle/install_ref.py while /root/reference is present; nothing here reads /root/reference).
# This pins the one restatement whose rounding the source text alone leaves open: is q*scale
# rounded to bf16 (my_fused_recurrent.py:59)?  It is not -- `scale` is a runtime fp32 scalar, so
# Triton promotes the product to fp32 -- which the bit-level comparison below establishes.
# ------------------------------------------------------------------------------------------------
def _ref_triton(variant):
    import os
    from oracle import install_ref
    rel = ("vllm_hymeta/models/my_fused_recurrent.py" if variant == "70b"
           else "vllm_hymeta/model_for_7B/my_fused_recurrent.py")
    if not os.path.isfile(os.path.join(install_ref.TARGET, rel)):
        pytest.skip("baseline/_ref (unmodified reference install) is not present on this box")
    try:
        return install_ref.load_ref_file(rel, f"_ref_mfr_{variant}")
    except Exception as e:
        pytest.skip(f"reference Triton module not importable here: {type(e).__name__}: {e}")


@pytest.mark.parametrize("variant", ["70b", "7b"])
def test_gla_decode_vs_reference_triton_kernel(ops, variant):
    """``my_fused_recurrent_fwd_kernel`` in the reference layout (one state per q-head, k/v/g
    repeat_kv'ed: Hs = H) against ``hymeta_gla_decode``: one step, then a 64-step chain through the
    in-place slot states, PAD slot included."""
    mod = _ref_triton(variant)
    B, H, Hkv = 5, 28, 4
    G = H // Hkv
    gen = torch.Generator().manual_seed(13)
    nslots = 9
    state0 = (torch.randn(nslots, H, D, D, generator=gen) * 0.1).bfloat16()
    slots = torch.tensor([7, 2, -1, 0, 5], dtype=torch.int32)
    live = slots >= 0
    st_ref, st_our, st_orc = state0.to(DEV), state0.to(DEV), state0.clone()
    steps = 64
    first = None
    for t in range(steps):
        q, k, v, g = ogla.synth_gla_inputs(B, H, Hkv, D, variant, seed=100 + t)
        rep = lambda x: x.repeat_interleave(G, dim=1).contiguous()
        k, v, g = rep(k), rep(v), rep(g)
        qd, kd, vd, gd = (x.to(DEV) for x in (q, k, v, g))
        o_ref = mod.my_fused_recurrent(qd.unsqueeze(2), kd.unsqueeze(2), vd.unsqueeze(2), gd.unsqueeze(2),
                                       st_ref, slots.to(DEV)).reshape(B, H, D)
        o_our = ops.gla_decode(qd, kd, vd, gd, st_our, slots.to(DEV))
        torch.cuda.synchronize()
        if t == 0:
            # one step from identical states: the oracle restatement vs the real kernel
            o_orc = ogla.gla_decode_step(q, k, v, g, st_orc, slots)
            s_ref, s_our = st_ref.cpu(), st_our.cpu()
            same_ref_orc = (s_ref[slots[live].long()] == st_orc[slots[live].long()]).float().mean().item()
            same_ref_our = (s_ref == s_our).float().mean().item()
            # exp() differs by an ulp between Triton's libdevice / torch CPU / ex2.approx, so the
            # state is compared at 1-bf16-ulp resolution and the bit-identical fraction is bounded
            assert_close(s_our, s_ref, "state after 1 step vs reference Triton kernel",
                         maxabs_frac=2 ** -8, rel_rms=2e-3)
            assert_close(st_orc, s_ref, "oracle state vs reference Triton kernel",
                         maxabs_frac=2 ** -8, rel_rms=2e-3)
            assert same_ref_our > 0.98 and same_ref_orc > 0.98, (same_ref_our, same_ref_orc)
            assert_close(o_our[live], o_ref[live], "output vs reference Triton kernel (1 step)")
            assert_close(o_orc[live], o_ref[live].cpu(), "oracle output vs reference Triton kernel")
            # q*scale is NOT rounded to bf16 by the reference kernel: an oracle variant that
            # rounds it is measurably further from the kernel's output than the fp32 one
            qr = (q.float() * D ** -0.5).bfloat16().float() / D ** -0.5
            o_rnd = ogla.gla_decode_step(qr.bfloat16(), k, v, g, state0.clone(), slots)
            e_fp32 = (o_orc[live].float() - o_ref[live].cpu().float()).abs().mean().item()
            e_rnd = (o_rnd[live].float() - o_ref[live].cpu().float()).abs().mean().item()
            first = (e_fp32, e_rnd)
    assert_close(st_our, st_ref, f"state after {steps} steps vs reference Triton kernel")
    assert_close(o_our[live], o_ref[live], f"output at step {steps} vs reference Triton kernel")
    # untouched slots (never addressed, and the PAD request) are bit-identical to the start
    idle = [s for s in range(nslots) if s not in slots.tolist()]
    assert torch.equal(st_our[idle].cpu(), state0[idle])
    assert first[0] <= first[1] + 1e-6, f"fp32 q*scale {first[0]:.3e} vs bf16-rounded {first[1]:.3e}"


def test_merge_vs_reference_torchscript(ops):
    """The reference's TorchScript ``_update_out_and_lse`` (fa_kernel_metatoken.py:110-122) run on
    the GPU against the oracle's restatement (already pinned on CPU by tests/golden/merge.npz)."""
    import os
    from oracle import install_ref
    rel = "vllm_hymeta/attention/fa_kernel_metatoken.py"
    if not os.path.isfile(os.path.join(install_ref.TARGET, rel)):
        pytest.skip("baseline/_ref is not present on this box")
    try:
        mod = install_ref.load_ref_file(rel, "_ref_fa_meta")
    except Exception as e:
        pytest.skip(f"reference module not importable here: {type(e).__name__}: {e}")
    gen = torch.Generator().manual_seed(3)
    o1 = torch.randn(1, 50, 4, D, generator=gen).bfloat16()
    o2 = torch.randn(1, 50, 4, D, generator=gen).bfloat16()
    l1 = torch.randn(1, 4, 50, generator=gen) * 3
    l2 = torch.randn(1, 4, 50, generator=gen) * 3
    ro, rl = mod._update_out_and_lse(o1.to(DEV), l1.to(DEV), o2.to(DEV), l2.to(DEV))
    oo, ol = oswa.merge_out_lse(o1, l1, o2, l2)
    assert_close(oo, ro, "merge out", maxabs_frac=2 ** -8, rel_rms=2e-3)
    assert torch.allclose(ol, rl.cpu(), atol=1e-5, rtol=1e-5)


# ------------------------------------------------------------------------------------------------
# vLLM's own CUDA ops (the installed 0.22 build; the reference pins 0.10.0): rotary_embedding
# (modeling_hymeta.py:230-231) and reshape_and_cache_flash (meta_attention.py:790-799)
# ------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def vllm_ops():
    try:
        import vllm._custom_ops as vops
        torch.ops._C.rotary_embedding          # noqa: B018  (raises if the extension is absent)
        torch.ops._C_cache_ops.reshape_and_cache_flash  # noqa: B018
        return vops
    except Exception as e:
        pytest.skip(f"vLLM custom ops not loadable here: {type(e).__name__}: {e}")


def test_kv_write_vs_vllm_reshape_and_cache_flash(ops, vllm_ops):
    T, Hkv, bs, nb = 777, 4, 256, 9
    gen = torch.Generator().manual_seed(17)
    key = torch.randn(T, Hkv, D, generator=gen).bfloat16().to(DEV)
    val = torch.randn(T, Hkv, D, generator=gen).bfloat16().to(DEV)
    cache0 = torch.randn(2, nb, bs, Hkv, D, generator=gen).bfloat16()
    slots = torch.randperm(nb * bs, generator=gen)[:T].to(torch.int64)
    slots[::13] = -1                                    # PAD entries are skipped by both
    ca, cb = cache0.to(DEV), cache0.to(DEV)
    one = torch.ones(1, dtype=torch.float32, device=DEV)
    vllm_ops.reshape_and_cache_flash(key, val, ca[0], ca[1], slots.to(DEV), "auto", one, one)
    ops.kv_write_paged(key, val, cb[0], cb[1], slots.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(ca.view(torch.int16), cb.view(torch.int16))


def test_rope_vs_vllm_rotary_embedding(ops, vllm_ops):
    """``hymeta_rope_kv_write_paged`` against the rotary op of the INSTALLED vLLM (0.22) on the same
    bf16 cos|sin table.  That build computes in fp32 and rounds once (tools/probe_rope_vllm.py), so
    ``fp32_math=True`` must reproduce it BIT FOR BIT; the default mode restates the per-operation
    bf16 rounding of vLLM 0.10.0's kernel (the version the reference pins, not installable here) and
    may differ from the 0.22 op by the two extra roundings only (<= 2 bf16 ulp)."""
    T, H, Hkv = 333, 28, 4
    gen = torch.Generator().manual_seed(19)
    q = torch.randn(T, H * D, generator=gen).bfloat16()
    k = torch.randn(T, Hkv * D, generator=gen).bfloat16()
    pos = torch.randint(0, 131072, (T,), generator=gen)
    table = ops.rope_cos_sin_cache(D, 131072, 1e6).to(DEV)
    qa, ka = q.to(DEV).clone(), k.to(DEV).clone()
    vllm_ops.rotary_embedding(pos.to(DEV), qa, ka, D, table, True)
    qb, kb = q.to(DEV).clone(), k.to(DEV).clone()
    ops.rope_kv_write_paged(qb.view(T, H, D), kb.view(T, Hkv, D), None, table, pos.to(DEV), fp32_math=True)
    qc, kc = q.to(DEV).clone(), k.to(DEV).clone()
    ops.rope_kv_write_paged(qc.view(T, H, D), kc.view(T, Hkv, D), None, table, pos.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(qa, qb) and torch.equal(ka, kb), "fp32 mode must equal the installed vLLM op"
    # the two extra roundings act on the products x*cos and y*sin, so the bound is relative to the
    # INPUT pair's magnitude (the rotated value itself may cancel to ~0): 2^-8 per rounded product
    # plus the final roundings of both results
    x = q.to(DEV).float().view(T, H, 2, D // 2)
    mag = (x[:, :, 0].abs() + x[:, :, 1].abs()).repeat(1, 1, 2).view(T, H * D)
    assert ((qa.float() - qc.float()).abs() <= 2.0 ** -7 * mag).all()
    assert (qa == qc).float().mean().item() > 0.5
