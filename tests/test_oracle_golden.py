"""Pin the CPU oracle against fixtures produced by the reference's own Python
(oracle/make_golden.py).  CPU only."""
import numpy as np
import torch

from oracle import gla, swa

T = torch.from_numpy


def close(a, b, atol, rtol=0.0):
    a, b = torch.as_tensor(a).float(), torch.as_tensor(b).float()
    assert a.shape == b.shape, (a.shape, b.shape)
    err = (a - b).abs().max().item()
    assert torch.allclose(a, b, atol=atol, rtol=rtol), f"max err {err}"


def test_gla_recurrence_vs_fla_naive(golden):
    g = golden("gla_naive.npz")
    for tag in ("7b", "70b"):
        q, k, v, gg, S0 = (T(g[f"{tag}_{n}"]) for n in ("q", "k", "v", "g", "S0"))
        o, S = gla.gla_recurrent(q, k, v, gg, S0)
        close(o, g[f"{tag}_o"], 2e-5, 1e-5)
        close(S, g[f"{tag}_S"], 2e-5, 1e-5)
        # compact (per-kv-head) initial state is the same thing when the group shares it
        o2, S2 = gla.gla_chunked(q, k, v, gg, S0, chunk=16)
        close(o2, g[f"{tag}_o"], 1e-4, 1e-4)
        close(S2, g[f"{tag}_S"], 1e-4, 1e-4)


def test_swa_dense_vs_reference_naive(golden):
    g = golden("swa_naive.npz")
    for tag in ("prefill", "suffix", "decode", "full"):
        W = int(g[f"{tag}_W"])
        # reference mask is (q_id - k_id) < W  ==  window_left = W - 1
        wl = W - 1 if W < (1 << 29) else -1
        q2 = T(g[f"{tag}_q2"]) if f"{tag}_q2" in g else None
        o, om = swa.swa_meta_dense(T(g[f"{tag}_q1"]), T(g[f"{tag}_k1"]), T(g[f"{tag}_v1"]), q2,
                                   T(g[f"{tag}_k2"]), T(g[f"{tag}_v2"]), wl)
        ref = g[f"{tag}_o"]                       # reference returns real-then-meta
        Lq = o.shape[0]
        close(o, ref[:Lq], 2e-6, 1e-5)
        if q2 is not None:
            close(om, ref[Lq:], 2e-6, 1e-5)
        two = swa.swa_meta_two_pass(T(g[f"{tag}_q1"]), T(g[f"{tag}_k1"]), T(g[f"{tag}_v1"]),
                                    T(g[f"{tag}_k2"]), T(g[f"{tag}_v2"]), wl)
        close(two, ref[:Lq], 5e-6, 1e-5)


def test_merge_vs_reference_torchscript(golden):
    g = golden("merge.npz")
    o, l = swa.merge_out_lse(T(g["out"]), T(g["lse"]), T(g["bout"]), T(g["blse"]))
    close(o, g["o"], 1e-6)
    close(l, g["l"], 1e-6)


def test_gate_7b_vs_hf_module(golden):
    g = golden("hf7b_gla_module.npz")
    H, Hkv, D = 4, 2, 128
    q_raw = T(g["q_raw"]).view(-1, H, D)
    k_raw = T(g["k_raw"]).view(-1, Hkv, D)
    v_raw = T(g["v_raw"]).view(-1, Hkv, D)
    gk_raw = T(g["gk_raw"]).view(-1, Hkv, D)
    q, k, gk = gla.gate_7b(q_raw, k_raw, gk_raw)
    # module tensors are head-first [H,T,D] and already repeat_kv'ed
    close(q.transpose(0, 1), g["q"], 1e-6)
    close(gla._expand_heads(k, H).transpose(0, 1), g["k"], 1e-6)
    close(gla._expand_heads(gk, H).transpose(0, 1), g["gk"], 1e-6)
    o, S = gla.gla_recurrent(q, k, v_raw, gk)
    close(o.transpose(0, 1), g["o"], 2e-5, 1e-5)
    close(S, g["h"], 2e-5, 1e-5)
    y = gla.rms_norm_head(o, T(g["g_norm_w"]), 1e-5).reshape(o.shape[0], -1) @ T(g["o_proj_w"]).T
    close(y, g["y"], 1e-4, 1e-4)


def test_gate_70b_vs_hf_module(golden):
    g = golden("hf70b_linear_module.npz")
    H, Hkv, D = 4, 2, 128
    for tag in ("first", "lb", "nolb"):
        q_raw = T(g[f"{tag}_q_raw"]).view(-1, H, D)
        k_raw = T(g[f"{tag}_k_raw"]).view(-1, Hkv, D)
        v_raw = T(g[f"{tag}_v_raw"]).view(-1, Hkv, D)
        lb = T(g[f"{tag}_lb"]) if f"{tag}_lb" in g else None
        # HF applies the gate after repeat_kv (lower bound is per q-head)
        k_rep = gla._expand_heads(k_raw, H)
        q, k, gg = gla.gate_70b(q_raw, k_rep, v_raw, lower_bound=lb,
                                layer_idx=int(g[f"{tag}_layer"]), num_heads=H, rule="hf")
        close(q.transpose(0, 1), g[f"{tag}_q"], 1e-6)
        close(k.transpose(0, 1), g[f"{tag}_k"], 2e-6)
        close(gg.transpose(0, 1), g[f"{tag}_g"], 1e-5, 1e-5)
        o, _ = gla.gla_recurrent(q, k, v_raw, gg)
        close(o.transpose(0, 1), g[f"{tag}_o"], 2e-5, 1e-5)


def test_rmsnorm_tp_branch_merge_vs_reference_class(golden):
    """f1: oracle.gla.rms_norm_tp / branch_norm_merge against the UNMODIFIED ``HymetaRMSNormTP``
    (modeling_hymeta.py:120-180) run at TP world 1, 2 and 4 by oracle/make_golden.py (ranks as
    threads, the all-reduce a summing barrier).  Bit-exact: same torch ops in the same order."""
    g = golden("hymeta_rmsnorm_tp.npz")
    x1, x2 = T(g["x_attn"]).bfloat16(), T(g["x_linear"]).bfloat16()
    w1, w2, eps = T(g["w1"]), T(g["w2"]), float(g["eps"])
    hidden = x1.shape[1]
    for W in (1, 2, 4):
        sl = [slice(r * hidden // W, (r + 1) * hidden // W) for r in range(W)]
        # what the summing all-reduce returns: the sum over ranks of each rank's local mean of squares
        tot1 = sum(x1[:, s].float().pow(2).mean(-1, keepdim=True, dtype=torch.float32) for s in sl)
        tot2 = sum(x2[:, s].float().pow(2).mean(-1, keepdim=True, dtype=torch.float32) for s in sl)
        got = torch.cat([gla.branch_norm_merge(x1[:, s], x2[:, s], w1[s], w2[s], eps, W,
                                               (lambda v, t=tot1: t, lambda v, t=tot2: t)) for s in sl], dim=1)
        assert torch.equal(got.float(), T(g[f"merged_tp{W}"])), W


def test_rope_vs_hf(golden):
    g = golden("rope.npz")
    pos = T(g["pos"])
    qe = swa.rope_neox(T(g["q"]).transpose(0, 1), pos)
    ke = swa.rope_neox(T(g["k"]).transpose(0, 1), pos)
    close(qe.transpose(0, 1), g["qe"], 1e-5)
    close(ke.transpose(0, 1), g["ke"], 1e-5)


def test_rope_table_form_vs_hf(golden):
    """The vLLM-side table formulation (cos|sin cache + x*cos - y*sin) equals the reference's HF
    rotary (golden, fp32) when evaluated in fp32; in bf16 it stays within bf16 rounding of it."""
    g = golden("rope.npz")
    pos = T(g["pos"]).long()
    table = swa.rope_cos_sin_cache(128, int(pos.max()) + 1, 1e6, dtype=torch.float32)
    for x, ref in ((g["q"], g["qe"]), (g["k"], g["ke"])):
        x = T(x).transpose(0, 1).contiguous()
        close(swa.rope_neox_table(x, pos, table).transpose(0, 1), ref, 1e-5)
        lo = swa.rope_neox_table(x.bfloat16(), pos, table.bfloat16()).float().transpose(0, 1)
        assert (lo - T(ref)).abs().max() <= 2.0 ** -6 * T(ref).abs().max()


def test_rope_kv_prologue_semantics():
    """Prologue restatement: meta rows captured, positions shifted by M, pages hold rotated k."""
    torch.manual_seed(3)
    Tn, M, H, Hkv, D, bs = 150, 16, 4, 2, 128, 16
    q, k, v = (torch.randn(Tn, h, D).bfloat16() for h in (H, Hkv, Hkv))
    table = swa.rope_cos_sin_cache(D, 4096, 10000.0)
    positions = torch.arange(Tn - M) + 7
    cache = torch.zeros(2, 12, bs, Hkv, D, dtype=torch.bfloat16)
    slots = torch.randperm(12 * bs)[: Tn - M]
    slots[3] = -1
    mk, mv = torch.zeros(M, Hkv, D).bfloat16(), torch.zeros(M, Hkv, D).bfloat16()
    qo, ko = swa.rope_kv_prologue(q, k, v, table, positions, M, cache[0], cache[1], slots, mk, mv, M)
    assert torch.equal(qo[:M], swa.rope_neox_table(q[:M], torch.arange(M), table))
    assert torch.equal(ko[M:], swa.rope_neox_table(k[M:], positions + M, table))
    assert torch.equal(mk, ko[:M]) and torch.equal(mv, v[:M])
    s = int(slots[0])
    assert torch.equal(cache[0, s // bs, s % bs], ko[M]) and torch.equal(cache[1, s // bs, s % bs], v[M])
    assert cache[0].flatten(0, 1)[torch.tensor(sorted(set(range(12 * bs)) - set(slots.tolist())))].abs().sum() == 0


def test_decode_step_matches_recurrence():
    """Faithful (bf16-rounding) decode steps track the fp32 recurrence (SURVEY 8c drift)."""
    torch.manual_seed(0)
    Tn, H, Hkv, D = 64, 4, 2, 128
    q, k, v, gg = gla.synth_gla_inputs(Tn, H, Hkv, D, "70b", seed=4)
    o_ref, S_ref = gla.gla_recurrent(q, k, v, gg)
    cache = torch.zeros(3, Hkv, D, D, dtype=torch.bfloat16)
    slot = torch.tensor([1], dtype=torch.int32)
    outs = [gla.gla_decode_step(q[t:t + 1], k[t:t + 1], v[t:t + 1], gg[t:t + 1], cache, slot)
            for t in range(Tn)]
    o = torch.cat(outs).float()
    assert (o - o_ref).abs().max() < 2 ** -6 * o_ref.abs().max()
    assert (cache[1].float() - S_ref[::2]).abs().max() < 0.05
    assert cache[0].abs().max() == 0 and cache[2].abs().max() == 0
    # PAD slot leaves everything untouched
    before = cache.clone()
    z = gla.gla_decode_step(q[:1], k[:1], v[:1], gg[:1], cache, torch.tensor([-1]))
    assert torch.equal(before, cache) and z.abs().max() == 0


def test_paged_decode_equals_dense_and_kv_write_bit_exact():
    torch.manual_seed(1)
    H, Hkv, D, M, bs = 4, 2, 64, 8, 16
    lens = [5, 37, 16, 64]
    nb = 16
    k_cache = torch.zeros(nb, bs, Hkv, D, dtype=torch.bfloat16)
    v_cache = torch.zeros_like(k_cache)
    perm = torch.randperm(nb).tolist()
    tables, slot_map, keys, vals = [], [], [], []
    it = iter(perm)
    for L in lens:
        row = [next(it) for _ in range((L + bs - 1) // bs)]
        tables.append(row + [0] * (4 - len(row)))
        slot_map += [row[p // bs] * bs + p % bs for p in range(L)]
        keys.append(torch.randn(L, Hkv, D).bfloat16())
        vals.append(torch.randn(L, Hkv, D).bfloat16())
    key, val = torch.cat(keys), torch.cat(vals)
    slot_map[3] = -1                                   # a padded token is skipped
    swa.kv_write_paged(key, val, k_cache, v_cache, slot_map)
    slot_map[3] = tables[0][0] * bs + 3
    swa.kv_write_paged(key[3:4], val[3:4], k_cache, v_cache, slot_map[3:4])
    off = 0
    for i, L in enumerate(lens):
        got = swa.gather_paged(k_cache, tables[i], 0, L)
        assert torch.equal(got, keys[i])
        off += L
    q = torch.randn(len(lens), H, D).bfloat16()
    k2, v2 = torch.randn(M, Hkv, D).bfloat16(), torch.randn(M, Hkv, D).bfloat16()
    for wl in (-1, 7, 20):
        out = swa.swa_meta_decode_paged(q, k_cache, v_cache, tables, lens, k2, v2, wl, out_dtype=torch.float32)
        for i, L in enumerate(lens):
            ref, _ = swa.swa_meta_dense(q[i:i + 1], keys[i], vals[i], None, k2, v2, wl)
            close(out[i], ref[0], 1e-6)
