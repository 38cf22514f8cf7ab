"""Torch-facing wrappers over the C-ABI kernels.  PyTorch is plumbing only: device
memory, the current stream and dtype/shape checks.  Every function enqueues on
``torch.cuda.current_stream()`` and never synchronises (CUDA-graph safe).
"""
from __future__ import annotations

import ctypes as C
import functools
from typing import Optional

import torch

from ._lib import check, lib

INIT_ZERO, INIT_SLOT, INIT_META = 0, 1, 2

_workspaces: dict = {}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _req(t: torch.Tensor, name: str, dtype=None):
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor (there is no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise ValueError(f"{name} must be {dtype}, got {t.dtype}")
    if t.stride(-1) != 1:
        raise ValueError(f"{name} must be contiguous in its last dimension")


# Optional per-launch timing for bench.py: between timing_begin() and timing_end() every wrapped op
# records a CUDA-event pair on the current stream around its launch (never during graph capture).
_TIMING = None


def timing_begin() -> None:
    global _TIMING
    _TIMING = []


def timing_end():
    """-> list of (op name, start event, end event) recorded since timing_begin()."""
    global _TIMING
    rec, _TIMING = _TIMING or [], None
    return rec


def _timed(name: str):
    def deco(fn):
        @functools.wraps(fn)
        def wrapper(*args, **kwargs):
            if _TIMING is None or torch.cuda.is_current_stream_capturing():
                return fn(*args, **kwargs)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*args, **kwargs)
            e1.record()
            _TIMING.append((name, e0, e1))
            return r
        return wrapper
    return deco


def launch_count() -> int:
    """Kernels launched by the library so far in this process."""
    return int(lib().hymeta_launch_count())


# ----------------------------------------------------------------------------------- a2
@_timed("gla_decode")
def gla_decode(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, g: torch.Tensor,
               state: torch.Tensor, slot_idx: torch.Tensor, scale: Optional[float] = None,
               out: Optional[torch.Tensor] = None, gnorm_weight: Optional[torch.Tensor] = None,
               gnorm_eps: float = 1e-6, gate_normalizer: Optional[float] = None) -> torch.Tensor:
    """q [B,H,D]; k,g [B,Hk,D]; v [B,Hv,D]; state [slots,Hs,D,D] bf16 (updated in place);
    slot_idx [B] int32 (-1 = skip).  Returns o [B,H,D] bf16.  gnorm_weight [D] bf16: the 7B
    per-head ``g_norm`` RMSNorm applied as the kernel's epilogue.  gate_normalizer: q, k and g
    are the RAW 7B projection rows (g = gk_proj output, bf16) and relu(q), relu(k),
    logsigmoid(gk) / gate_normalizer run inside the kernel."""
    B, H, D = q.shape
    Hk, Hv, Hs = k.shape[1], v.shape[1], state.shape[1]
    if g.shape[1] != Hk:
        raise ValueError("g must have as many heads as k")
    for n, t in (("q", q), ("k", k), ("v", v), ("state", state)):
        _req(t, n, torch.bfloat16)
    _req(g, "g")
    if g.dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("g must be float32 or bfloat16")
    _req(slot_idx, "slot_idx", torch.int32)
    if state.stride(2) != D or state.shape[2] != D or state.shape[3] != D:
        raise ValueError("state must be [slots, Hs, D, D] with contiguous [D, D] tiles")
    if out is None:
        out = torch.empty(B, H, D, dtype=q.dtype, device=q.device)
    if scale is None:
        scale = D ** -0.5
    if gnorm_weight is not None:
        _req(gnorm_weight, "gnorm_weight", torch.bfloat16)
        if gnorm_weight.numel() != D:
            raise ValueError("gnorm_weight must hold head_dim elements")
    if gate_normalizer is not None and (gate_normalizer <= 0 or g.dtype != torch.bfloat16):
        raise ValueError("the fused 7B gate needs a positive normalizer and bf16 gk rows")
    check(lib().hymeta_gla_decode_gate7b(
        _ptr(q), _ptr(k), _ptr(v), _ptr(g), _ptr(out), _ptr(state), _ptr(slot_idx),
        B, H, Hk, Hv, Hs, D, q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0), v.stride(1),
        g.stride(0), g.stride(1), out.stride(0), out.stride(1), state.stride(0), state.stride(1),
        int(g.dtype == torch.float32), float(scale), _ptr(gnorm_weight), float(gnorm_eps),
        float(gate_normalizer or 0.0), _stream()), "hymeta_gla_decode")
    return out


# ----------------------------------------------------------------------------------- a3
@_timed("gla_prefill")
def gla_prefill_varlen(q, k, v, g, state, cu_seqlens, slot_idx, init_mode,
                       meta_state: Optional[torch.Tensor] = None, scale: Optional[float] = None,
                       write_state: bool = True, max_seqlen: Optional[int] = None,
                       out: Optional[torch.Tensor] = None, seg_len: Optional[int] = None,
                       gnorm_weight: Optional[torch.Tensor] = None,
                       gnorm_eps: float = 1e-6, relu_qk: bool = False,
                       gate_table: Optional[torch.Tensor] = None,
                       sumsq_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q [T,H,D]; k,g [T,Hk,D]; v [T,Hv,D]; state [slots,Hs,D,D]; cu_seqlens [N+1], slot_idx [N],
    init_mode [N] int32 device tensors.  Returns o [T,H,D] bf16; writes final states.
    gnorm_weight [D] bf16: the 7B per-head ``g_norm`` RMSNorm applied as the kernel's epilogue.
    relu_qk: q and k are raw projection rows, relu is applied in the kernel.
    gate_table (``gla_gate_7b_table``): g holds the RAW bf16 gk rows and the kernel maps them through
    the table -- no separate gate pass.  seg_len: None = automatic (sequence-parallel segments for few
    long chains, ``GLA_SEG_BALANCED`` when there are more chains than SMs, else one chain per CTA), 0 =
    one chain per CTA, > 0 = segment length.  sumsq_out [T,H] fp32: the epilogue also writes each bf16
    output row's sum of squares (f1: RMS-norm partials for ``branch_norm_merge``)."""
    T, H, D = q.shape
    Hk, Hv, Hs = k.shape[1], v.shape[1], state.shape[1]
    if g.shape[1] != Hk:
        raise ValueError("g must have as many heads as k")
    N = cu_seqlens.numel() - 1
    for n, t in (("q", q), ("k", k), ("v", v), ("state", state)):
        _req(t, n, torch.bfloat16)
    _req(g, "g")
    for n, t in (("cu_seqlens", cu_seqlens), ("slot_idx", slot_idx), ("init_mode", init_mode)):
        _req(t, n, torch.int32)
    if meta_state is not None:
        _req(meta_state, "meta_state", torch.bfloat16)
    if out is None:
        out = torch.empty(T, H, D, dtype=q.dtype, device=q.device)
    if scale is None:
        scale = D ** -0.5
    msl = int(max_seqlen if max_seqlen is not None else T)
    if seg_len is None:
        seg_len = gla_prefill_seg_len(N * Hs, msl, q.device)
        if seg_len == 0 and gla_prefill_balanced_pays(N * Hs, msl, _sm_count(q.device)):
            seg_len = GLA_SEG_BALANCED
    ws, ws_bytes = None, 0
    if seg_len == GLA_SEG_BALANCED:
        ws_bytes = lib().hymeta_gla_prefill_balanced_workspace_bytes(N, Hs)
        if ws_bytes:
            ws = _gla_workspace(q.device, ws_bytes)
        else:
            seg_len = 0
    elif seg_len:
        ws_bytes = lib().hymeta_gla_prefill_workspace_bytes(N, Hs, msl, seg_len)
        if ws_bytes:
            ws = _gla_workspace(q.device, ws_bytes)
    if gnorm_weight is not None:
        _req(gnorm_weight, "gnorm_weight", torch.bfloat16)
        if gnorm_weight.numel() != D:
            raise ValueError("gnorm_weight must hold head_dim elements")
    if gate_table is not None:
        _req(gate_table, "gate_table", torch.bfloat16)
        if gate_table.numel() != 65536 or not gate_table.is_contiguous() or g.dtype != torch.bfloat16:
            raise ValueError("gate_table must be a contiguous [65536] bf16 tensor and g the raw bf16 gate rows")
    check(lib().hymeta_gla_prefill_varlen_fused(
        _ptr(q), _ptr(k), _ptr(v), _ptr(g), _ptr(out), _ptr(state), _ptr(meta_state),
        _ptr(cu_seqlens), _ptr(slot_idx), _ptr(init_mode), N, T, H, Hk, Hv, Hs, D,
        q.stride(0), q.stride(1), k.stride(0), k.stride(1), v.stride(0), v.stride(1),
        g.stride(0), g.stride(1), out.stride(0), out.stride(1), state.stride(0), state.stride(1),
        0 if meta_state is None else meta_state.stride(0),
        int(g.dtype == torch.float32), float(scale), int(write_state), msl,
        int(seg_len or 0), _ptr(ws), ws_bytes, _ptr(gnorm_weight), float(gnorm_eps), int(bool(relu_qk)),
        _ptr(gate_table), _ptr(_sumsq_buf(sumsq_out, T, H)), _stream()), "hymeta_gla_prefill_varlen_seg")
    return out


def _sumsq_buf(t: Optional[torch.Tensor], rows: int, heads: int) -> Optional[torch.Tensor]:
    if t is not None:
        _req(t, "sumsq_out", torch.float32)
        if t.shape != (rows, heads) or not t.is_contiguous():
            raise ValueError(f"sumsq_out must be a contiguous [{rows}, {heads}] fp32 tensor")
    return t


_SM_COUNT = {}
_WS = {}
GLA_SEG_BALANCED = -1      # HYMETA_GLA_SEG_BALANCED: equal step ranges per SM, chains cut where they straddle


def _gla_workspace(device, nbytes: int) -> torch.Tensor:
    """Grow-only fp32 scratch per device (the C ABI never allocates).  One buffer per device:
    segmented prefills are single-stream calls.  It must already be large enough when a CUDA graph
    is captured (a buffer allocated from the graph's private pool must not be cached for eager use)."""
    key = (device.type, device.index)
    w = _WS.get(key)
    if w is None or w.numel() * 4 < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("gla_prefill workspace must be sized by an eager call before CUDA-graph capture")
        w = torch.empty((nbytes + 3) // 4, dtype=torch.float32, device=device)
        _WS[key] = w
    return w


def _sm_count(device=None) -> int:
    dev = torch.cuda.current_device() if device is None or device.index is None else device.index
    sms = _SM_COUNT.get(dev)
    if sms is None:
        sms = _SM_COUNT[dev] = torch.cuda.get_device_properties(dev).multi_processor_count
    return sms


def gla_prefill_balanced_pays(chains: int, max_seqlen: int, sms: int) -> bool:
    """Whether the balanced work lists (equal step ranges per SM, chains cut where they straddle) beat
    one chain per CTA.  They remove the partial last wave -- ceil(w) / w - 1 with w = chains / SMs
    waves -- and pay ~4 steps of pipeline drain + restart per piece plus ~6 % for the static split
    (a hardware-scheduled grid evens out SM-to-SM speed differences, fixed lists cannot).  Measured
    on B200 (tools/bench_gla_balanced.py; chains x steps): 160 x 512 -41 %, 256 x 512 -9 %,
    400 x 256 -5 %, 1200 x 128 -5 %, 1024 x 64 +13 %, 4000 x 16 +25 %.  The host only knows the
    longest sequence, so the estimate takes every chain at that length."""
    if chains <= sms or max_seqlen <= 0:
        return False
    steps = -(-max_seqlen // 16)
    w = chains / sms
    gain = -(-chains // sms) / w - 1.0
    return gain > 4.0 / steps + 0.06 + 0.02


def gla_prefill_seg_len(chains: int, max_seqlen: int, device=None, min_seg: int = 256) -> int:
    """Segment length for the sequence-parallel prefill, or 0 for the single-pass form.  One CTA per
    SM walks one chain, so the launch is quantised in waves of ``sms`` CTAs: candidates are the
    segment counts that just fill 1, 2, 4 or 8 waves (chains * n_segs <= w * sms).  Measured on B200
    (tools/sweep_gla_segs.py) the three-pass form costs ~1.5x its ideal w * T / n_segs (the summary
    pass re-reads k/v/g), so it is taken only when that still beats ceil(chains / sms) full-length
    waves -- never for whole waves of full sequences (config 2), always for few long chains."""
    sms = _sm_count(device)
    if chains <= 0 or max_seqlen < 2 * min_seg:
        return 0
    best_cost, best_ns = float(-(-chains // sms)), 1          # single pass: full-length waves
    for w in (1, 2, 4, 8):
        ns = min((w * sms) // chains, max_seqlen // min_seg)
        if ns < 2:
            continue
        # + a small per-segment term (workspace state round trip, pipeline ramp): among fills that
        # model the same, the sweep favours the fewer, longer segments
        cost = -(-chains * ns // sms) * 1.5 / ns + 0.002 * ns
        if cost < best_cost - 1e-9:
            best_cost, best_ns = cost, ns
    if best_ns <= 1:
        return 0
    seg = -(-max_seqlen // best_ns)
    return -(-seg // 16) * 16


# ----------------------------------------------------------------------------------- a4
@_timed("gla_gate_70b")
def gla_gate_70b(q: torch.Tensor, k: torch.Tensor, lower_bound: Optional[torch.Tensor] = None,
                 clamp_max: float = 0.95, inplace_q: bool = False):
    """70B feature map + gate in one launch.  q [T,H,D], k [T,Hk,D] bf16 (strided views of the
    qkv projection are fine).  ``lower_bound``: fp32 [Ho,D] rank-local slice or None.
    Returns (relu(q) [T,H,D] bf16, k' [T,Ho,D] bf16, g [T,Ho,D] fp32), Ho = Hk without a lower
    bound, else lower_bound.shape[0] (H in the reference)."""
    T, H, D = q.shape
    Hk = k.shape[1]
    _req(q, "q", torch.bfloat16)
    _req(k, "k", torch.bfloat16)
    Ho = Hk
    if lower_bound is not None:
        _req(lower_bound, "lower_bound", torch.float32)
        if not lower_bound.is_contiguous() or lower_bound.shape[-1] != D:
            raise ValueError("lower_bound must be a contiguous fp32 [Ho, D] tensor")
        Ho = lower_bound.numel() // D
    q_out = q if inplace_q else torch.empty(T, H, D, dtype=q.dtype, device=q.device)
    k_out = torch.empty(T, Ho, D, dtype=k.dtype, device=k.device)
    g_out = torch.empty(T, Ho, D, dtype=torch.float32, device=k.device)
    check(lib().hymeta_gla_gate_70b(
        _ptr(q), _ptr(k), _ptr(q_out), _ptr(k_out), _ptr(g_out), _ptr(lower_bound), T, H, Hk, Ho, D,
        q.stride(0), q.stride(1), k.stride(0), k.stride(1), q_out.stride(0), q_out.stride(1),
        k_out.stride(0), k_out.stride(1), g_out.stride(0), g_out.stride(1), float(clamp_max),
        _stream()), "hymeta_gla_gate_70b")
    return q_out, k_out, g_out


@_timed("gla_gate_7b")
def gla_gate_7b(q: torch.Tensor, k: torch.Tensor, gk: torch.Tensor, normalizer: float = 16.0,
                inplace: bool = False):
    """7B feature map + gate in one launch.  q [T,H,D], k, gk [T,Hk,D] bf16.
    Returns (relu(q), relu(k), logsigmoid(gk)/normalizer) all bf16."""
    T, H, D = q.shape
    Hk = k.shape[1]
    for n, t in (("q", q), ("k", k), ("gk", gk)):
        _req(t, n, torch.bfloat16)
    q_out = q if inplace else torch.empty(T, H, D, dtype=q.dtype, device=q.device)
    k_out = k if inplace else torch.empty(T, Hk, D, dtype=k.dtype, device=k.device)
    g_out = gk if inplace else torch.empty(T, Hk, D, dtype=gk.dtype, device=gk.device)
    check(lib().hymeta_gla_gate_7b(
        _ptr(q), _ptr(k), _ptr(gk), _ptr(q_out), _ptr(k_out), _ptr(g_out), T, H, Hk, D,
        q.stride(0), q.stride(1), k.stride(0), k.stride(1), gk.stride(0), gk.stride(1),
        q_out.stride(0), q_out.stride(1), k_out.stride(0), k_out.stride(1),
        g_out.stride(0), g_out.stride(1), float(normalizer), _stream()), "hymeta_gla_gate_7b")
    return q_out, k_out, g_out


@_timed("gla_gate_7b")
def gla_gate_7b_g(gk: torch.Tensor, normalizer: float = 16.0, out: Optional[torch.Tensor] = None):
    """Gate half of the 7B pre-processing only: g = bf16(bf16(logsigmoid(gk)) / normalizer), [T,Hk,D]
    bf16.  relu(q) / relu(k) are applied by ``gla_prefill_varlen(relu_qk=True)`` on the raw rows."""
    T, Hk, D = gk.shape
    _req(gk, "gk", torch.bfloat16)
    g_out = out if out is not None else torch.empty(T, Hk, D, dtype=gk.dtype, device=gk.device)
    check(lib().hymeta_gla_gate_7b_g(_ptr(gk), _ptr(g_out), T, Hk, D, gk.stride(0), gk.stride(1),
                                     g_out.stride(0), g_out.stride(1), float(normalizer), _stream()),
          "hymeta_gla_gate_7b_g")
    return g_out


_GATE_TABLES = {}


def gla_gate_7b_table(normalizer: float = 16.0, device=None) -> torch.Tensor:
    """[65536] bf16: ``gla_gate_7b_g`` applied to every bf16 bit pattern -- the 7B gate
    bf16(bf16(logsigmoid(gk)) / normalizer) is a pure function of its 16-bit input, so the GLA prefill
    kernel's gate warps look it up (``gla_prefill_varlen(gate_table=...)``) instead of a separate pass
    over the activations.  Built once per (device, normalizer) with the gate kernel itself, hence
    bit-identical to it."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    key = (device.index, float(normalizer))
    t = _GATE_TABLES.get(key)
    if t is None:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("build the gate table with an eager call before CUDA-graph capture")
        bits = torch.arange(65536, dtype=torch.int32, device=device).to(torch.int16).view(torch.bfloat16)
        t = _GATE_TABLES[key] = gla_gate_7b_g(bits.view(512, 1, 128), normalizer).view(-1).contiguous()
    return t


# ----------------------------------------------------------------------------------- a9
@_timed("kv_write")
def kv_write_paged(key, value, k_cache, v_cache, slot_mapping) -> None:
    """key,value [T,Hkv,D]; caches [nb,bs,Hkv,D]; slot_mapping [T] int64 (<0 = skip)."""
    T, Hkv, D = key.shape
    for n, t in (("key", key), ("value", value), ("k_cache", k_cache), ("v_cache", v_cache)):
        _req(t, n, torch.bfloat16)
    _req(slot_mapping, "slot_mapping", torch.int64)
    if slot_mapping.numel() < T:
        raise ValueError("slot_mapping shorter than the number of tokens")
    check(lib().hymeta_kv_write_paged(
        _ptr(key), _ptr(value), _ptr(k_cache), _ptr(v_cache), _ptr(slot_mapping), T, Hkv, D,
        k_cache.shape[1], key.stride(0), key.stride(1), value.stride(0), value.stride(1),
        k_cache.stride(0), k_cache.stride(1), k_cache.stride(2),
        v_cache.stride(0), v_cache.stride(1), v_cache.stride(2), _stream()), "hymeta_kv_write_paged")


# ----------------------------------------------------------------------------------- a10
def advance_step(num_seqs: int, num_queries: int, block_size: int, seq_lens: torch.Tensor,
                 slot_mapping: torch.Tensor, block_tables: torch.Tensor,
                 input_tokens: Optional[torch.Tensor] = None,
                 sampled_token_ids: Optional[torch.Tensor] = None,
                 input_positions: Optional[torch.Tensor] = None) -> None:
    """Device-side metadata advance of multi-step decode (vLLM ``advance_step_flashattn``):
    seq_lens [num_seqs] int32 += 1 for the live rows, slot_mapping [num_seqs] int64 = slot of the new
    last position through block_tables [num_seqs, max_blocks] int32; padded rows get slot -1."""
    _req(seq_lens, "seq_lens", torch.int32)
    _req(slot_mapping, "slot_mapping", torch.int64)
    _req(block_tables, "block_tables", torch.int32)
    for n, t in (("input_tokens", input_tokens), ("sampled_token_ids", sampled_token_ids),
                 ("input_positions", input_positions)):
        if t is not None:
            _req(t, n, torch.int64)
            if t.numel() < num_queries:
                raise ValueError(f"{n} shorter than num_queries")
    if seq_lens.numel() < num_seqs or slot_mapping.numel() < num_seqs or block_tables.shape[0] < num_seqs:
        raise ValueError("seq_lens / slot_mapping / block_tables shorter than num_seqs")
    check(lib().hymeta_advance_step(int(num_seqs), int(num_queries), int(block_size), _ptr(input_tokens),
                                    _ptr(sampled_token_ids), _ptr(input_positions), _ptr(seq_lens),
                                    _ptr(slot_mapping), _ptr(block_tables), block_tables.stride(0),
                                    _stream()), "hymeta_advance_step")


# ----------------------------------------------------------------------------------- f3
def rope_cos_sin_cache(head_dim: int, max_position: int, base: float, device=None,
                       dtype=torch.bfloat16) -> torch.Tensor:
    """The table HymetaRotaryEmbedding builds (modeling_hymeta.py:205-218): fp32
    [cos(t*inv_freq) | sin(t*inv_freq)] cast to ``dtype``; [max_position, head_dim]."""
    inv_freq = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float) / head_dim))
    freqs = torch.einsum("i,j -> ij", torch.arange(max_position, dtype=torch.float), inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype).to(device)


@_timed("rope_kv_write")
def rope_kv_write_paged(q, k, v, cos_sin_cache, positions, pos_offset=0, k_cache=None, v_cache=None,
                        slot_mapping=None, meta_k=None, meta_v=None, n_meta=0, q_out=None,
                        k_out=None, fp32_math: bool = False):
    """NeoX RoPE of q [T,H,D] and k [T,Hkv,D] (in place unless q_out/k_out are given), rotated
    k and v [T,Hkv,D] written to their pages, the first ``n_meta`` tokens captured into
    meta_k/meta_v instead -- one launch.  Returns (q_out, k_out).
    fp32_math=False: bf16 rounding after every product and sum (the vLLM 0.10.0 op the reference
    pins); True: fp32 math, one rounding (bit-identical to the installed vLLM 0.22 op)."""
    T, H, D = q.shape
    Hkv = k.shape[1]
    q_out = q if q_out is None else q_out
    k_out = k if k_out is None else k_out
    if v is None:                       # rotate only (HymetaRotaryEmbedding.forward)
        if n_meta > 0 or (k_cache is not None and k_cache.numel() > 0):
            raise ValueError("v is required when pages or meta rows are written")
        v = k                           # never dereferenced: no destination for value rows
    for n, t in (("q", q), ("k", k), ("v", v), ("q_out", q_out), ("k_out", k_out),
                 ("cos_sin_cache", cos_sin_cache)):
        _req(t, n, torch.bfloat16)
    if k.shape[0] != T or v.shape != k.shape or q_out.shape != q.shape or k_out.shape != k.shape:
        raise ValueError("q/k/v/q_out/k_out shapes disagree")
    if cos_sin_cache.dim() != 2 or cos_sin_cache.shape[1] != D or not cos_sin_cache.is_contiguous():
        raise ValueError("cos_sin_cache must be a contiguous [max_position, head_dim] table")
    if T > n_meta:
        _req(positions, "positions", torch.int64)
        if positions.numel() < T - n_meta:
            raise ValueError("positions shorter than the number of non-meta tokens")
    paged = k_cache is not None and k_cache.numel() > 0
    if paged:
        _req(k_cache, "k_cache", torch.bfloat16)
        _req(v_cache, "v_cache", torch.bfloat16)
        _req(slot_mapping, "slot_mapping", torch.int64)
        if slot_mapping.numel() < T - n_meta:
            raise ValueError("slot_mapping shorter than the number of non-meta tokens")
        kc = (k_cache.stride(0), k_cache.stride(1), k_cache.stride(2))
        vc = (v_cache.stride(0), v_cache.stride(1), v_cache.stride(2))
        bs = k_cache.shape[1]
    else:
        kc = vc = (0, 0, 0)
        bs = 0
    if n_meta > 0:
        _req(meta_k, "meta_k", torch.bfloat16)
        _req(meta_v, "meta_v", torch.bfloat16)
        if meta_k.shape[0] < n_meta or meta_v.shape[0] < n_meta:
            raise ValueError("meta_k/meta_v hold fewer rows than n_meta")
        mk, mv = (meta_k.stride(0), meta_k.stride(1)), (meta_v.stride(0), meta_v.stride(1))
    else:
        mk = mv = (0, 0)
    check(lib().hymeta_rope_kv_write_paged(
        _ptr(q), _ptr(k), _ptr(v), _ptr(q_out), _ptr(k_out), _ptr(cos_sin_cache),
        cos_sin_cache.shape[0], _ptr(positions) if T > n_meta else None, int(pos_offset),
        _ptr(k_cache) if paged else None, _ptr(v_cache) if paged else None,
        _ptr(slot_mapping) if paged else None,
        _ptr(meta_k) if n_meta > 0 else None, _ptr(meta_v) if n_meta > 0 else None,
        n_meta, T, H, Hkv, D, bs, q.stride(0), q.stride(1), k.stride(0), k.stride(1),
        v.stride(0), v.stride(1), q_out.stride(0), q_out.stride(1), k_out.stride(0),
        k_out.stride(1), *kc, *vc, *mk, *mv, int(bool(fp32_math)), _stream()),
        "hymeta_rope_kv_write_paged")
    return q_out, k_out


# ----------------------------------------------------------------------------------- f1
def branch_sumsq(x_attn: torch.Tensor, x_linear: torch.Tensor,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Per-token mean of x^2 over the rank's hidden slice, both branches: [2, T] fp32."""
    T, Hd = x_attn.shape
    _req(x_attn, "x_attn", torch.bfloat16)
    _req(x_linear, "x_linear", torch.bfloat16)
    if x_linear.shape != x_attn.shape:
        raise ValueError("the two branches must have the same shape")
    if out is None:
        out = torch.empty(2, T, dtype=torch.float32, device=x_attn.device)
    check(lib().hymeta_branch_sumsq(_ptr(x_attn), _ptr(x_linear), _ptr(out), T, Hd,
                                    x_attn.stride(0), x_linear.stride(0), _stream()),
          "hymeta_branch_sumsq")
    return out


class BranchMergeExchange:
    """Peer-mapped exchange buffers of the fused tensor-parallel branch merge (f1): one buffer per rank
    of the TP group, allocated as torch symmetric memory (CUDA peer mapping over NVLink) and handed
    to ``hymeta_branch_norm_merge_tp`` as raw addresses.  Create it once per group (collective call:
    every rank of ``group`` must construct it with the same ``t_cap``); each ``branch_norm_merge(...,
    exchange=ex)`` call advances the epoch, so every rank must make the same sequence of calls.
    ``device_epoch`` (default): the epoch lives in device memory and the kernel advances it itself, so
    the call can be captured in a CUDA graph and replayed (every rank replays the same graphs in the
    same order); False: the host counts and passes the epoch by value (eager launches only)."""

    def __init__(self, group, t_cap: int, device=None, device_epoch: bool = True):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        if not 2 <= self.world <= 8:
            raise ValueError("the fused exchange serves tensor-parallel groups of 2..8 ranks")
        self.t_cap = int(t_cap)
        nbytes = lib().hymeta_branch_merge_tp_buffer_bytes(self.world, self.t_cap)
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, self.group)
        self.ptrs = (C.c_uint64 * self.world)(*[int(self.handle.buffer_ptrs[r]) for r in range(self.world)])
        torch.cuda.synchronize(device)
        dist.barrier(self.group)        # every buffer is zeroed before any peer may write into it
        self.epoch = 0
        # {last epoch, CTAs done}: read and advanced by the kernel (hymeta_branch_norm_merge_tp_graph)
        self.epoch_state = torch.zeros(2, dtype=torch.int32, device=device) if device_epoch else None

    def next_epoch(self) -> int:
        self.epoch = self.epoch % 0x7fffffff + 1      # never 0 (the zero fill), +1 per call
        return self.epoch


@_timed("branch_norm_merge")
def branch_norm_merge(x_attn: torch.Tensor, x_linear: torch.Tensor, w_attn: torch.Tensor,
                      w_linear: torch.Tensor, eps: float, tp_world: int = 1,
                      means: Optional[torch.Tensor] = None, group=None,
                      out: Optional[torch.Tensor] = None,
                      exchange: Optional[BranchMergeExchange] = None,
                      sumsq: Optional[tuple] = None) -> torch.Tensor:
    """``((norm1(out_attn) + norm2(out_linear)) / 2).to(bfloat16)`` of the 70B layer
    (modeling_hymeta.py:909-919, HymetaRMSNormTP :151-173).  x_* [T, hidden/tp] bf16, w_* the
    rank's fp32 weight slice.  tp_world > 1 with ``exchange``: ONE kernel computes the local means,
    exchanges them through NVLink peer memory and merges with the rows still in registers.  Without
    it: ``branch_sumsq`` + ONE NCCL all-reduce of 8*T bytes over ``group`` + the merge kernel (pass
    ``means`` to supply an already reduced buffer).  ``sumsq`` = (sumsq_attn, sumsq_linear), the
    [T, heads_local] fp32 partials the attention kernels' epilogues emitted (``sumsq_out=`` of the
    prefill ops): with ``exchange`` the merge then reads x exactly once."""
    T, Hd = x_attn.shape
    for n, t in (("x_attn", x_attn), ("x_linear", x_linear)):
        _req(t, n, torch.bfloat16)
    for n, t in (("w_attn", w_attn), ("w_linear", w_linear)):
        _req(t, n, torch.float32)
        if t.numel() != Hd:
            raise ValueError(f"{n} must hold hidden/tp = {Hd} elements")
    if x_linear.shape != x_attn.shape:
        raise ValueError("the two branches must have the same shape")
    if tp_world > 1 and exchange is not None and means is None:
        if exchange.world != tp_world or T > exchange.t_cap:
            raise ValueError("exchange was built for another group size or a smaller token capacity")
        if out is None:
            out = torch.empty(T, Hd, dtype=torch.bfloat16, device=x_attn.device)
        sa = sl = None
        heads = 0
        if sumsq is not None:
            sa, sl = sumsq
            heads = sa.shape[1]
            for t in (sa, sl):
                _req(t, "sumsq", torch.float32)
                if t.shape != (T, heads) or not t.is_contiguous():
                    raise ValueError("sumsq tensors must be contiguous [T, heads_local] fp32")
        if exchange.epoch_state is not None:
            check(lib().hymeta_branch_norm_merge_tp_graph(
                _ptr(x_attn), _ptr(x_linear), _ptr(w_attn), _ptr(w_linear), _ptr(out), T, Hd, tp_world,
                exchange.rank, exchange.ptrs, exchange.t_cap, _ptr(exchange.epoch_state), float(eps),
                x_attn.stride(0), x_linear.stride(0), out.stride(0), _ptr(sa), _ptr(sl), heads, _stream()),
                "hymeta_branch_norm_merge_tp_graph")
            return out
        check(lib().hymeta_branch_norm_merge_tp_sumsq(
            _ptr(x_attn), _ptr(x_linear), _ptr(w_attn), _ptr(w_linear), _ptr(out), T, Hd, tp_world,
            exchange.rank, exchange.ptrs, exchange.t_cap, exchange.next_epoch(), float(eps),
            x_attn.stride(0), x_linear.stride(0), out.stride(0), _ptr(sa), _ptr(sl), heads, _stream()),
            "hymeta_branch_norm_merge_tp")
        return out
    if tp_world > 1 and means is None:
        import torch.distributed as dist
        means = branch_sumsq(x_attn, x_linear)
        dist.all_reduce(means, group=group)
    if means is not None:
        _req(means, "means", torch.float32)
        if means.numel() != 2 * T or not means.is_contiguous():
            raise ValueError("means must be a contiguous [2, T] fp32 buffer")
    if out is None:
        out = torch.empty(T, Hd, dtype=torch.bfloat16, device=x_attn.device)
    check(lib().hymeta_branch_norm_merge(
        _ptr(x_attn), _ptr(x_linear), _ptr(means), _ptr(w_attn), _ptr(w_linear), _ptr(out), T, Hd,
        int(tp_world), float(eps), x_attn.stride(0), x_linear.stride(0), out.stride(0), _stream()),
        "hymeta_branch_norm_merge")
    return out


# ----------------------------------------------------------------------------------- a8
def _workspace(nbytes: int, device) -> torch.Tensor:
    key = device.index        # one workspace per device: calls on one device are stream-ordered by the caller
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("workspace must be sized before CUDA-graph capture "
                               "(run one eager step first)")
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


_tickets: dict = {}


def _ticket_buffer(nbytes: int, device) -> torch.Tensor:
    """Zero-initialised int32 tickets of the fused split-K combine (the kernel restores zero)."""
    t = _tickets.get(device.index)
    if t is None or t.numel() * 4 < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("ticket buffer must be sized before CUDA-graph capture "
                               "(run one eager step first)")
        t = torch.zeros(max((nbytes + 3) // 4, 4096), dtype=torch.int32, device=device)
        _tickets[device.index] = t
    return t


@_timed("swa_decode")
def swa_meta_decode_paged(q, k_cache, v_cache, block_tables, cache_seqlens, meta_k, meta_v,
                          window_left: int, scale: Optional[float] = None,
                          out: Optional[torch.Tensor] = None,
                          new_k: Optional[torch.Tensor] = None, new_v: Optional[torch.Tensor] = None,
                          fused_combine: bool = False,
                          new_slot_mapping: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q [B,H,D]; caches [nb,bs,Hkv,D]; block_tables [B,max_blocks] int32; cache_seqlens [B]
    int32; meta_k/meta_v [M,Hkv,D] or None.  Returns out [B,H,D].
    new_k/new_v [B,Hkv,D]: K/V of the token at position cache_seqlens[b]-1; the kernel writes them to
    their page (reshape_and_cache_flash folded in) and attends to them -- no separate kv_write_paged.
    new_slot_mapping [B] int64: destination slots of the new tokens (< 0 = padded row, not stored);
    without it the page of position cache_seqlens[b]-1 is taken from the block table and padded
    rows MUST carry cache_seqlens = 0.
    fused_combine: the split-K partials are merged by the last split CTA to finish instead of a
    second launch (bit-identical; measured no faster on B200, so off by default).  With a single
    split per (request, kv-head) -- the plan whenever the batch fills the GPU -- the split kernel
    writes the final rows itself and there is no combine at all."""
    B, H, D = q.shape
    Hkv, bs = k_cache.shape[2], k_cache.shape[1]
    M = 0 if meta_k is None else meta_k.shape[0]
    for n, t in (("q", q), ("k_cache", k_cache), ("v_cache", v_cache)):
        _req(t, n, torch.bfloat16)
    _req(block_tables, "block_tables", torch.int32)
    _req(cache_seqlens, "cache_seqlens", torch.int32)
    if M:
        _req(meta_k, "meta_k", torch.bfloat16)
        _req(meta_v, "meta_v", torch.bfloat16)
    if (new_k is None) != (new_v is None):
        raise ValueError("new_k and new_v must be given together")
    if new_k is not None:
        _req(new_k, "new_k", torch.bfloat16)
        _req(new_v, "new_v", torch.bfloat16)
        if new_k.shape != (B, Hkv, D) or new_v.shape != (B, Hkv, D):
            raise ValueError("new_k/new_v must be [B, Hkv, D]")
        if new_slot_mapping is not None:
            _req(new_slot_mapping, "new_slot_mapping", torch.int64)
            if new_slot_mapping.numel() < B:
                raise ValueError("new_slot_mapping shorter than the batch")
    elif new_slot_mapping is not None:
        raise ValueError("new_slot_mapping needs new_k/new_v")
    if out is None:
        out = torch.empty(B, H, D, dtype=q.dtype, device=q.device)
    if scale is None:
        scale = D ** -0.5
    max_blocks = block_tables.shape[1]
    nbytes = lib().hymeta_swa_decode_workspace_bytes(B, H, Hkv, D, max_blocks, bs, window_left, M)
    ws = _workspace(nbytes, q.device)
    tickets = _ticket_buffer(lib().hymeta_swa_decode_ticket_bytes(B, H, Hkv), q.device) \
        if fused_combine and B > 0 else None
    check(lib().hymeta_swa_meta_decode_paged_fused(
        _ptr(q), _ptr(k_cache), _ptr(v_cache), _ptr(block_tables), _ptr(cache_seqlens),
        _ptr(meta_k), _ptr(meta_v), _ptr(out), _ptr(ws), ws.numel(),
        B, H, Hkv, D, M, bs, k_cache.shape[0], max_blocks, block_tables.stride(0), int(window_left),
        float(scale),
        q.stride(0), q.stride(1), out.stride(0), out.stride(1),
        k_cache.stride(0), k_cache.stride(1), k_cache.stride(2),
        v_cache.stride(0), v_cache.stride(1), v_cache.stride(2),
        meta_k.stride(0) if M else 0, meta_k.stride(1) if M else 0,
        meta_v.stride(0) if M else 0, meta_v.stride(1) if M else 0,
        _ptr(new_k), _ptr(new_v),
        new_k.stride(0) if new_k is not None else 0, new_k.stride(1) if new_k is not None else 0,
        new_v.stride(0) if new_v is not None else 0, new_v.stride(1) if new_v is not None else 0,
        _ptr(new_slot_mapping), _ptr(tickets), _stream()),
        "hymeta_swa_meta_decode_paged_fused")
    return out


# ----------------------------------------------------------------------------------- a7
@_timed("swa_prefill")
def swa_meta_prefill_varlen(q1, k1, v1, q2, meta_k, meta_v, cu_seqlens_q, cu_seqlens_k,
                            max_seqlen_q: int, window_left: int, scale: Optional[float] = None,
                            block_tables: Optional[torch.Tensor] = None,
                            out: Optional[torch.Tensor] = None,
                            sumsq_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q1 [Tq,H,D]; k1,v1 [Tk,Hkv,D] (or paged caches [nb,bs,Hkv,D] with block_tables);
    q2 [M,H,D] or None; meta_k/meta_v [M,Hkv,D] or None.  Returns [(M if q2)+Tq, H, D], meta
    rows first.  sumsq_out [rows,H] fp32: the epilogue also writes each bf16 output row's sum of
    squares (f1: RMS-norm partials for ``branch_norm_merge``)."""
    Tq, H, D = q1.shape
    paged = block_tables is not None
    Hkv = k1.shape[-2]
    M = 0 if meta_k is None else meta_k.shape[0]
    for n, t in (("q1", q1), ("k1", k1), ("v1", v1)):
        _req(t, n, torch.bfloat16)
    _req(cu_seqlens_q, "cu_seqlens_q", torch.int32)
    _req(cu_seqlens_k, "cu_seqlens_k", torch.int32)
    if q2 is not None:
        _req(q2, "q2", torch.bfloat16)
        if M == 0 or q2.shape[0] != M:
            raise ValueError("q2 requires meta keys with the same number of rows")
    if M:
        _req(meta_k, "meta_k", torch.bfloat16)
        _req(meta_v, "meta_v", torch.bfloat16)
    N = cu_seqlens_q.numel() - 1
    rows = Tq + (M if q2 is not None else 0)
    if out is None:
        out = torch.empty(rows, H, D, dtype=q1.dtype, device=q1.device)
    if scale is None:
        scale = D ** -0.5
    if paged:
        _req(block_tables, "block_tables", torch.int32)
        if k1.dim() != 4 or v1.shape != k1.shape:
            raise ValueError("paged prefill: k1/v1 must be the [num_blocks, block_size, Hkv, D] caches")
        if block_tables.dim() != 2 or block_tables.shape[0] < cu_seqlens_q.numel() - 1:
            raise ValueError("block_tables must be [num_seqs, max_blocks]")
        Tk, bs = k1.shape[0] * k1.shape[1], k1.shape[1]     # rows of the cache = num_blocks * block_size
        k_st, k_sh, v_st, v_sh = k1.stride(1), k1.stride(2), v1.stride(1), v1.stride(2)
        kc_sblk, vc_sblk, bt_stride = k1.stride(0), v1.stride(0), block_tables.stride(0)
    else:
        Tk, bs = k1.shape[0], 0
        k_st, k_sh, v_st, v_sh = k1.stride(0), k1.stride(1), v1.stride(0), v1.stride(1)
        kc_sblk = vc_sblk = bt_stride = 0
    check(lib().hymeta_swa_meta_prefill_varlen_sumsq(
        _ptr(q1), _ptr(k1), _ptr(v1), _ptr(q2), _ptr(meta_k), _ptr(meta_v), _ptr(out),
        _ptr(cu_seqlens_q), _ptr(cu_seqlens_k), _ptr(block_tables), bt_stride, bs,
        N, Tq, Tk, H, Hkv, D, M, int(max_seqlen_q), int(window_left), float(scale),
        q1.stride(0), q1.stride(1), k_st, k_sh, v_st, v_sh,
        q2.stride(0) if q2 is not None else 0, q2.stride(1) if q2 is not None else 0,
        meta_k.stride(0) if M else 0, meta_k.stride(1) if M else 0,
        meta_v.stride(0) if M else 0, meta_v.stride(1) if M else 0,
        out.stride(0), out.stride(1), kc_sblk, vc_sblk, _ptr(_sumsq_buf(sumsq_out, rows, H)), _stream()),
        "hymeta_swa_meta_prefill_varlen")
    return out


# ----------------------------------------------------------------------------------- a5
def state_zero_slots(state: torch.Tensor, slots: torch.Tensor) -> None:
    """state [L, slots, Hs, D, D] bf16; zero state[:, slots[i]] for every i (slots int32, <0 skipped)."""
    _req(state, "state", torch.bfloat16)
    _req(slots, "slots", torch.int32)
    L = state.shape[0]
    slot_elems = state[0, 0].numel()
    if not state[0, 0].is_contiguous():
        raise ValueError("each (layer, slot) state block must be contiguous")
    check(lib().hymeta_state_zero_slots(_ptr(state), _ptr(slots), slots.numel(), L,
                                        state.stride(0), state.stride(1), slot_elems, _stream()),
          "hymeta_state_zero_slots")


def state_copy_slot(state: torch.Tensor, from_slot: int, to_slot: int) -> None:
    _req(state, "state", torch.bfloat16)
    L = state.shape[0]
    slot_elems = state[0, 0].numel()
    check(lib().hymeta_state_copy_slot(_ptr(state), int(from_slot), int(to_slot), L,
                                       state.stride(0), state.stride(1), slot_elems, _stream()),
          "hymeta_state_copy_slot")


def selftest_umma() -> float:
    import ctypes as C
    err = C.c_float(0.0)
    rc = lib().hymeta_selftest_umma(C.addressof(err))
    if rc != 0:
        raise RuntimeError(f"tcgen05/TMA self test failed at case {rc} (max err {err.value})")
    return err.value
