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
