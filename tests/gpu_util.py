"""Shared helpers for the -m gpu parity tests (CUDA path vs CPU oracle)."""
import torch

DEV = "cuda"

# Tolerances: bf16 outputs vs the fp32 oracle
MAXABS_FRAC = 2.0 ** -7      # max-abs error <= 2^-7 * max|ref|
REL_RMS = 5e-3               # relative RMS error (test_vllm/allclose.py protocol)


def assert_close(got: torch.Tensor, ref: torch.Tensor, what: str, maxabs_frac=MAXABS_FRAC,
                 rel_rms=REL_RMS):
    got, ref = got.detach().float().cpu(), ref.detach().float().cpu()
    assert got.shape == ref.shape, (what, got.shape, ref.shape)
    assert torch.isfinite(got).all(), f"{what}: non-finite output"
    scale = max(ref.abs().max().item(), 1e-6)
    err = (got - ref).abs().max().item()
    rms = ((got - ref).pow(2).mean().sqrt() / max(ref.pow(2).mean().sqrt().item(), 1e-12)).item()
    assert err <= maxabs_frac * scale, f"{what}: max-abs {err:.4g} > {maxabs_frac:.3g}*{scale:.4g}"
    assert rms <= rel_rms, f"{what}: rel-RMS {rms:.4g} > {rel_rms}"
    return err, rms


def make_paged(lens, Hkv, D, block_size, seed, extra_blocks=3, max_blocks=None):
    """Random paged KV cache: physical blocks are a seeded permutation so gathers are
    non-trivial.  Returns (k_cache, v_cache, block_tables[int32], keys list, vals list)."""
    gen = torch.Generator().manual_seed(seed)
    nblk = [(L + block_size - 1) // block_size for L in lens]
    nb = sum(nblk) + extra_blocks
    k_cache = torch.randn(nb, block_size, Hkv, D, generator=gen).bfloat16()   # garbage elsewhere
    v_cache = torch.randn(nb, block_size, Hkv, D, generator=gen).bfloat16()
    perm = torch.randperm(nb, generator=gen).tolist()
    mb = max_blocks or max(max(nblk), 1)
    tables = torch.zeros(len(lens), mb, dtype=torch.int32)
    keys, vals = [], []
    it = iter(perm)
    for i, L in enumerate(lens):
        row = [next(it) for _ in range(nblk[i])]
        tables[i, :len(row)] = torch.tensor(row, dtype=torch.int32)
        kk = torch.randn(L, Hkv, D, generator=gen).bfloat16()
        vv = torch.randn(L, Hkv, D, generator=gen).bfloat16()
        for p in range(L):
            k_cache[row[p // block_size], p % block_size] = kk[p]
            v_cache[row[p // block_size], p % block_size] = vv[p]
        keys.append(kk)
        vals.append(vv)
    return k_cache, v_cache, tables, keys, vals
