"""-m gpu parity: GLA chunked prefill (tcgen05) through the C-ABI vs the CPU oracle recurrence."""
import pytest
import torch

from gpu_util import DEV, assert_close
from oracle import gla as ogla

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import __graft_entry__ as ge
    ge.build()
    from vllm_hymeta_b200 import ops as o
    return o


def run_case(ops, lens, H, Hkv, model, layout, seed, modes=None, slots=None, write_state=True,
             g_override=None, seg_len=0):
    D = 128
    T = sum(lens)
    q, k, v, g = ogla.synth_gla_inputs(T, H, Hkv, D, model, seed=seed)
    if g_override is not None:
        g = torch.full_like(g, g_override)
    Hs = Hkv if layout == "compact" else H
    if layout == "reference":
        k, v, g = (x.repeat_interleave(H // Hkv, 1).contiguous() for x in (k, v, g))
    N = len(lens)
    gen = torch.Generator().manual_seed(seed + 100)
    nslots = N + 2
    state = (torch.randn(nslots, Hs, D, D, generator=gen) * 0.2).bfloat16()
    meta = (torch.randn(Hs, D, D, generator=gen) * 0.2).bfloat16()
    slots = slots if slots is not None else torch.randperm(nslots, generator=gen)[:N].tolist()
    modes = modes if modes is not None else [i % 3 for i in range(N)]
    cu = [0]
    for L in lens:
        cu.append(cu[-1] + L)
    ref_state = state.clone()
    ref_o = ogla.gla_prefill_varlen(q, k, v, g, ref_state, cu, slots, modes, meta,
                                    write_state=write_state, out_dtype=torch.float32)
    d_state = state.to(DEV)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    o = ops.gla_prefill_varlen(q.to(DEV), k.to(DEV), v.to(DEV), g.to(DEV), d_state, i32(cu),
                               i32(slots), i32(modes), meta.to(DEV), write_state=write_state,
                               max_seqlen=max(lens), seg_len=seg_len)
    torch.cuda.synchronize()
    assert_close(o, ref_o, f"gla prefill o {lens} {model} {layout}")
    assert_close(d_state, ref_state, f"gla prefill state {lens} {model} {layout}")
    used = set(s for s in slots if s >= 0) if write_state else set()
    for s in range(nslots):
        if s not in used:
            assert torch.equal(d_state[s].cpu(), state[s]), f"slot {s} modified"
    return o, d_state


@pytest.mark.parametrize("model", ["7b", "70b"])
def test_single_sequence(ops, model):
    run_case(ops, [100], 28, 4, model, "compact", 1, modes=[0])


@pytest.mark.parametrize("model", ["7b", "70b"])
def test_ragged_slots_and_init_modes(ops, model):
    run_case(ops, [1, 15, 16, 17, 200, 33, 64], 28, 4, model, "compact", 2)


def test_reference_and_mixed_layouts(ops):
    run_case(ops, [40, 70], 8, 2, "70b", "reference", 3)      # one state per q-head, repeated k/v/g
    run_case(ops, [40, 70], 8, 2, "7b", "mixed", 4)           # one state per q-head, un-repeated k/v/g


def test_tp4_shard_geometry(ops):
    run_case(ops, [300], 7, 1, "70b", "compact", 5, modes=[2])


def test_strong_decay_at_clamp_limit(ops):
    """70B worst case: k clamped at 0.95 -> g = log(0.05) = -2.996 on every channel and step."""
    run_case(ops, [130], 14, 2, "70b", "compact", 6, modes=[1], g_override=-2.9957)


def test_no_decay(ops):
    run_case(ops, [90], 14, 2, "7b", "compact", 7, modes=[1], g_override=0.0)


def test_write_state_false_and_pad_slot(ops):
    run_case(ops, [50, 20], 28, 4, "70b", "compact", 8, write_state=False)
    run_case(ops, [50, 20], 28, 4, "70b", "compact", 9, slots=[-1, 1], modes=[0, 1])


def test_long_sequence_and_prefill_decode_consistency(ops):
    """T=4096 vs the fp32 oracle, and the size-independent property
    prefill(T) == prefill(T-8) followed by 8 recurrent decode steps (same final state / outputs)."""
    H, Hkv, D, T = 28, 4, 128, 4096
    q, k, v, g = ogla.synth_gla_inputs(T, H, Hkv, D, "70b", seed=11)
    ref_o, ref_S = ogla.gla_chunked(q, k, v, g, chunk=16)
    dq, dk, dv, dg = (x.to(DEV) for x in (q, k, v, g))
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    # the prefix property is bit-exact whenever both runs cut the sequence at the same places:
    # single pass (seg_len 0) and a fixed segment length
    for seg_len in (0, 512):
        st = torch.zeros(2, Hkv, D, D, dtype=torch.bfloat16, device=DEV)
        o = ops.gla_prefill_varlen(dq, dk, dv, dg, st, i32([0, T]), i32([0]), i32([0]), seg_len=seg_len)
        assert_close(o, ref_o, "long o")
        assert_close(st[0], ref_S[::H // Hkv], "long state")
        o2 = ops.gla_prefill_varlen(dq[:T - 8], dk[:T - 8], dv[:T - 8], dg[:T - 8], st,
                                    i32([0, T - 8]), i32([1]), i32([0]), seg_len=seg_len)
        tail = [ops.gla_decode(dq[t:t + 1], dk[t:t + 1], dv[t:t + 1], dg[t:t + 1], st, i32([1]))
                for t in range(T - 8, T)]
        assert torch.equal(o2, o[:T - 8])
        assert_close(torch.cat(tail), ref_o[T - 8:], "decode tail")
        assert_close(st[1], st[0].float(), "state after prefill+decode", maxabs_frac=2.0 ** -6, rel_rms=1e-2)


@pytest.mark.parametrize("model", ["7b", "70b"])
def test_segmented_matches_oracle(ops, model):
    """Sequence-parallel form (summaries -> scan -> outputs): ragged lengths around the segment size,
    every init mode, a PAD slot and a sequence shorter than one segment."""
    run_case(ops, [700, 130, 1000, 256, 257, 1], 28, 4, model, "compact", 11, seg_len=256,
             modes=[1, 2, 0, 1, 2, 1], slots=[3, 0, -1, 5, 1, 2])


def test_segmented_equals_single_pass(ops):
    o1, s1 = run_case(ops, [2048], 7, 1, "70b", "compact", 12, modes=[2], seg_len=0)
    o2, s2 = run_case(ops, [2048], 7, 1, "70b", "compact", 12, modes=[2], seg_len=128)
    o3, s3 = run_case(ops, [2048], 7, 1, "70b", "compact", 12, modes=[2], seg_len=None)   # auto
    for o, s in ((o2, s2), (o3, s3)):
        assert (o.float() - o1.float()).abs().max().item() <= 2 ** -6 * o1.float().abs().max().item()
        assert (s.float() - s1.float()).abs().max().item() <= 2 ** -6 * s1.float().abs().max().item()


def test_seg_len_heuristic():
    from vllm_hymeta_b200 import ops as o
    assert o.gla_prefill_seg_len(256, 8192) == 0            # full waves: single pass
    assert o.gla_prefill_seg_len(4, 300) == 0               # too short to cut
    s = o.gla_prefill_seg_len(1, 32768)
    assert s > 0 and s % 16 == 0 and 32768 // s >= 64
    # balanced work lists: only where the partial last wave outweighs the per-piece restart
    assert o.gla_prefill_balanced_pays(256, 8192, 148)      # config 2: 1.73 waves
    assert o.gla_prefill_balanced_pays(160, 8192, 148)
    assert not o.gla_prefill_balanced_pays(148, 8192, 148)  # one full wave
    assert not o.gla_prefill_balanced_pays(296, 8192, 148)  # two full waves
    assert not o.gla_prefill_balanced_pays(4000, 256, 148)  # many short chains: the grid wins


# ------------------------------------------------------------------- balanced single-pass form
def _balanced_vs_single(ops, lens, H, Hkv, Hs, model, seed, with_sumsq=False):
    """seg_len = GLA_SEG_BALANCED must reproduce the single-pass launch bit for bit: the chains are
    only re-distributed over the SMs, a cut hands the fp32 state (and its pending decay) over."""
    D, T, N = 128, sum(lens), len(lens)
    q, k, v, g = ogla.synth_gla_inputs(T, H, Hkv, D, model, seed=seed)
    if Hs != Hkv:
        k, v, g = (x.repeat_interleave(Hs // Hkv, 1).contiguous() for x in (k, v, g))
    gen = torch.Generator().manual_seed(seed + 5)
    nslots = N + 2
    state = (torch.randn(nslots, Hs, D, D, generator=gen) * 0.2).bfloat16()
    meta = (torch.randn(Hs, D, D, generator=gen) * 0.2).bfloat16().to(DEV)
    slots = torch.randperm(nslots, generator=gen)[:N].tolist()
    slots[3] = -1                                            # a PAD request
    modes = [i % 3 for i in range(N)]
    cu = [0]
    for L in lens:
        cu.append(cu[-1] + L)
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=DEV)
    dq, dk, dv, dg = (x.to(DEV) for x in (q, k, v, g))
    res = []
    for seg_len in (0, ops.GLA_SEG_BALANCED):
        st = state.to(DEV)
        ss = torch.full((T, H), -1.0, device=DEV) if with_sumsq else None
        o = torch.full((T, H, D), float("nan"), device=DEV, dtype=torch.bfloat16)
        ops.gla_prefill_varlen(dq, dk, dv, dg, st, i32(cu), i32(slots), i32(modes), meta,
                               max_seqlen=max(lens), seg_len=seg_len, out=o, sumsq_out=ss)
        torch.cuda.synchronize()
        res.append((o, st, ss))
    (o0, s0, q0), (o1, s1, q1) = res
    assert torch.isfinite(o0.float()).all()
    assert torch.equal(o1, o0), f"outputs differ: {(o1.float() - o0.float()).abs().max().item()}"
    assert torch.equal(s1, s0), "final states differ"
    if with_sumsq:
        assert torch.equal(q1, q0)
    return o0


def test_balanced_equals_single_pass_more_chains_than_sms(ops):
    gen = torch.Generator().manual_seed(3)
    # 90 requests x 4 state heads = 360 chains on 148 SMs; ragged tails, a 1-token and two empty requests
    lens = torch.randint(48, 330, (90,), generator=gen).tolist()
    lens[0], lens[7], lens[20], lens[89] = 1, 0, 16, 0
    _balanced_vs_single(ops, lens, 28, 4, 4, "7b", 21, with_sumsq=True)
    _balanced_vs_single(ops, lens, 28, 4, 4, "70b", 22)


def test_balanced_one_state_per_q_head(ops):
    lens = [130, 64, 257, 300, 90, 111, 200, 31]              # x 28 heads = 224 chains, G = 1
    _balanced_vs_single(ops, lens, 28, 4, 28, "70b", 23)


def test_balanced_falls_back_when_a_chain_exceeds_an_sm_share(ops):
    lens = [40] * 60 + [4000]                                 # 244 chains, one far longer than total / 148
    _balanced_vs_single(ops, lens, 28, 4, 4, "7b", 24)
    _balanced_vs_single(ops, [100, 50, 60, 70], 28, 4, 4, "7b", 25)   # fewer chains than SMs: plain grid


def test_balanced_many_short_requests(ops):
    """More requests than plan threads (N > 1024) and far more chains than SMs: every SM's list holds
    dozens of pieces, most of them whole chains of a few steps, some empty."""
    gen = torch.Generator().manual_seed(5)
    lens = torch.randint(0, 70, (1500,), generator=gen).tolist()
    _balanced_vs_single(ops, lens, 8, 2, 2, "70b", 26)
