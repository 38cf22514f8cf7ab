"""-m gpu parity: fused SWA prologue (RoPE + paged-KV write + meta capture) through
the C ABI vs the oracle's step-by-step restatement.  Everything here is bit-exact: the kernel
reproduces the per-operation bf16 rounding of the reference's rotary op and the page write is a
copy."""
import pytest
import torch

from oracle import swa as oswa
from gpu_util import DEV

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from vllm_hymeta_b200 import ops as o
    return o


CASES = [
    # (T, n_meta, H, Hkv, block_size, pos_offset, paged)
    (777, 0, 28, 4, 256, 0, True),          # 7B prefill chunk, no meta tokens
    (128 + 900, 128, 28, 4, 256, 128, True),  # 70B first pass: 128 meta tokens captured
    (64, 0, 28, 4, 16, 128, True),          # decode batch (one token per request), 70B offset
    (300, 16, 7, 1, 64, 16, True),          # TP=4 shard
    (200, 0, 28, 4, 256, 0, False),         # profiling run: no pages, rotate only
    (128, 128, 28, 4, 256, 128, True),      # meta tokens only
]


@pytest.mark.parametrize("case", CASES)
def test_rope_kv_write_bit_exact(ops, case):
    T, M, H, Hkv, bs, off, paged = case
    D = 128
    gen = torch.Generator().manual_seed(T + M)
    qkv = torch.randn(T, (H + 2 * Hkv) * D, generator=gen).bfloat16()
    qkv[1, :4] = torch.tensor([1e-39, -1e-39, 0.0, 3e38]).bfloat16()    # denormals / huge survive
    split = lambda x: (x[:, :H * D].view(T, H, D), x[:, H * D:(H + Hkv) * D].view(T, Hkv, D),
                       x[:, (H + Hkv) * D:].view(T, Hkv, D))
    q, k, v = split(qkv)
    max_pos = 9000
    table = oswa.rope_cos_sin_cache(D, max_pos, 10000.0)
    positions = torch.randint(0, max_pos - off, (T - M,), generator=gen, dtype=torch.int64)
    nb = (T + bs - 1) // bs + 3
    cache = torch.randn(2, nb, bs, Hkv, D, generator=gen).bfloat16()
    slots = torch.randperm(nb * bs, generator=gen)[: T - M].to(torch.int64)
    if T - M > 6:
        slots[5] = -1
    mk = torch.randn(max(M, 1), Hkv, D, generator=gen).bfloat16()
    mv = torch.randn(max(M, 1), Hkv, D, generator=gen).bfloat16()
    ref_cache, ref_mk, ref_mv = cache.clone(), mk.clone(), mv.clone()
    rq, rk = oswa.rope_kv_prologue(q, k, v, table, positions, off,
                                   ref_cache[0] if paged else None, ref_cache[1] if paged else None,
                                   slots, ref_mk, ref_mv, M)
    d_qkv, d_cache, d_mk, d_mv = qkv.to(DEV), cache.to(DEV), mk.to(DEV), mv.to(DEV)
    dq, dk, dv = split(d_qkv)
    v_before = dv.clone()
    d_table = ops.rope_cos_sin_cache(D, max_pos, 10000.0, device=DEV)
    assert torch.equal(d_table.cpu(), table)
    oq, ok = ops.rope_kv_write_paged(
        dq, dk, dv, d_table, positions.to(DEV), off,
        d_cache[0] if paged else None, d_cache[1] if paged else None,
        slots.to(DEV) if paged else None, d_mk if M else None, d_mv if M else None, M)
    torch.cuda.synchronize()
    assert oq.data_ptr() == dq.data_ptr() and ok.data_ptr() == dk.data_ptr()    # in place
    bits = lambda x: x.cpu().contiguous().view(torch.int16)
    assert torch.equal(bits(oq), bits(rq)), "rotated q differs"
    assert torch.equal(bits(ok), bits(rk)), "rotated k differs"
    assert torch.equal(dv, v_before), "v must not be touched"
    assert torch.equal(bits(d_cache), bits(ref_cache)), "pages differ"
    if M:
        assert torch.equal(bits(d_mk), bits(ref_mk)) and torch.equal(bits(d_mv), bits(ref_mv))


def test_rope_kv_out_of_place_and_errors(ops):
    T, H, Hkv, D = 40, 28, 4, 128
    gen = torch.Generator().manual_seed(5)
    q = torch.randn(T, H, D, generator=gen).bfloat16().to(DEV)
    k = torch.randn(T, Hkv, D, generator=gen).bfloat16().to(DEV)
    v = torch.randn(T, Hkv, D, generator=gen).bfloat16().to(DEV)
    table = ops.rope_cos_sin_cache(D, 512, 1e6, device=DEV)
    pos = torch.arange(T, device=DEV)
    q0, k0 = q.clone(), k.clone()
    qo, ko = torch.empty_like(q), torch.empty_like(k)
    ops.rope_kv_write_paged(q, k, v, table, pos, 0, q_out=qo, k_out=ko)
    torch.cuda.synchronize()
    assert torch.equal(q, q0) and torch.equal(k, k0)
    ref = oswa.rope_neox_table(q0.cpu(), pos.cpu(), table.cpu())
    assert torch.equal(qo.cpu().view(torch.int16), ref.view(torch.int16))
    ops.rope_kv_write_paged(q[:0], k[:0], v[:0], table, pos[:0], 0)            # empty: no-op
    with pytest.raises(ValueError):
        ops.rope_kv_write_paged(q.cpu(), k, v, table, pos, 0)                  # no CPU path
    with pytest.raises(ValueError):
        ops.rope_kv_write_paged(q, k, v, table, pos.int(), 0)                  # positions int64
    with pytest.raises(ValueError):
        ops.rope_kv_write_paged(q, k, v, table[:, :64], pos, 0)                # rotary_dim != D
