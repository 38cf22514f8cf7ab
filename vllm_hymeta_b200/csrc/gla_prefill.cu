// GLA chunked prefill on tcgen05 / TMEM (a1/a3), varlen over cu_seqlens with slot
// states -- replaces the reference's per-request Python loop around fused_chunk_gla, its two host
// syncs per request and the 7x repeat_kv materialisation (modeling_hymeta.py:441-549).
//
// One CTA (one per SM, all 512 TMEM columns) walks one (sequence, state-head[, segment]) chain in
// steps of C = 16 tokens.  The G = H/Hs q-heads sharing the state are stacked along the MMA M
// dimension: row = head*16 + token (G <= 8 -> M = 128).  With b_t = cumsum(g) inside the step and
// the per-channel reference r = b_7 (mid-step, so every exponent is bounded by 8 steps of decay):
//     q~_t = q_t * scale * exp(b_t - r)      k~_s = k_s * exp(r - b_s)
//     A    = q~ k~^T  (causal, s <= t)               tcgen05.mma  M128 N16  K128   (K-major x K-major)
//     O    = q~ X  +  P v      (P = bf16 causal A)    tcgen05.mma  M128 N128 K128 (SS) + K16 (TS, P in TMEM)
//     Y    = X + k~^T v                               mma.sync m16n8k16, accumulators = the fp32 master
//     X'   = exp(b_15 - r + r') (.) Y                 rescaled in registers
// where X = exp(r) (.) S is the state expressed in the step's reference frame.  The fp32 master lives
// in the registers of the state warpgroup (with it in TMEM every step would move 128 KB through
// tcgen05.ld / st for the per-row decay); its bf16 copy -- the MN-major B operand of the next step's
// q~ X -- is written to shared memory by the pass that rescales it.
//
// Warp roles (512 threads, setmaxnreg 80 / 208 / 112), all hand-offs through mbarriers:
//   warps 0-1   gates: thread = 2 channels; raw g / k rows arrive by TMA (2-slot ring), v by TMA
//               straight into its stage; builds k~ (relu folded in for the 7B layer), the q-side
//               factor table scale*exp(b_t - r) and the per-channel rescale factors, 3-stage ring
//   warp 2      MMA issuer (one lane);  warp 3 idle
//   warps 4-7   state: mma.sync Y = X + k~^T v on ldmatrix fragments, rescale, bf16 X to smem,
//               final state / segment summary store
//   warps 8-11  q-tile builder (q prefetched one tile ahead, relu folded in) / P converter /
//               O epilogue (+ g_norm) for even steps (thread = row for TMEM); output rows staged per
//               warp in smem and stored as 64-byte row segments
//   warps 12-15 same for odd steps
#include "umma.cuh"

namespace hymeta {

constexpr int GC = 16;                          // tokens per step
#ifndef GP_KS_
#define GP_KS_ 3
#endif
#ifndef GP_GN_UNROLL
#define GP_GN_UNROLL 4   // g_norm pass 1: 4 chunks in flight (A/B on B200: 1 -> 2.55, 2 -> 2.50, 4 -> 2.47, 8 -> 2.54 ms)
#endif
#define GP_STR_(x) #x
#define GP_STR(x) GP_STR_(x)
#ifndef GP_EXP
#define GP_EXP 0   // timing / bisection experiments of tools/trace_gla_prefill.py (bit mask); 0 in the product build
#endif
constexpr int GP_KS = GP_KS_;                   // gate ring stages
constexpr int GP_SQ = 0;                        // q~  2 x [2][128][64] bf16            64 KB
constexpr int GP_SS = 65536;                    // X   2 x [2][128 dk][64 dv] bf16      64 KB
constexpr int GP_ST = 131072;                   // gate ring
constexpr int GP_ST_K = 0;                      //   k~ [2][16][64] bf16                 4 KB
constexpr int GP_ST_V = 4096;                   //   v  [2][16][64] bf16                 4 KB
constexpr int GP_ST_F = 8192;                   //   scale*exp(b_t - r) [16][128] f32    8 KB
constexpr int GP_ST_D = 16384;                  //   exp(b_15 - r) [128], exp(r) [128]   1 KB
constexpr int GP_STAGE = 17408;
constexpr int GP_RAW = GP_ST + GP_KS * GP_STAGE;   // raw g / k ring of the gate warps (cp.async)
constexpr int GP_RS = 2;                        //   stages
constexpr int GP_RAW_K = 8192;                  //   g [16][64 thr] 4|8 B (8 KB), k [16][64 thr] 4 B (4 KB)
constexpr int GP_RAW_STAGE = 12288;
constexpr int GP_OST = GP_RAW + GP_RS * GP_RAW_STAGE;   // output staging: 8 q-tile warps x 2 KB
constexpr int GP_BAR = GP_OST + 8 * 2048;
constexpr int GP_GW = GP_BAR + 256;             // g_norm weight [128] bf16 (7B epilogue)
constexpr int GP_SMEM = GP_GW + 256;
constexpr int GP_THREADS = 512;
constexpr int GP_WS_ELEMS = kD * kD + kD;      // per (chain, segment): state [128][128] + decay [128]

// One unit of a CTA's work: a run of whole steps of one (sequence, state-head) chain.  A launch over
// the (Hs, N, segments) grid gives every CTA the one piece its block index names; the balanced
// single-pass form (gla_prefill_plan_kernel below) gives each CTA a list.
struct __align__(16) GlaPiece {
  int32_t tok_begin, len;   // first token row and token count
  int32_t hs, in_kind;      // state head; start state: 0 zero, 1 slot in_idx, 2 meta state, 3 fp32 buffer in_idx
  int32_t in_idx, out_kind; // end state: 0 dropped, 1 slot out_idx (bf16), 2 fp32 buffer out_idx
  int32_t out_idx, pad;
};
constexpr int GP_PLAN_HDR = 4;                  // plan header ints: {cut mode, ctas, ticket, steps per CTA}


struct GlaPrefillParams {
  const __nv_bfloat16 *q, *k, *v;
  const void* g;
  __nv_bfloat16 *o, *state;
  const __nv_bfloat16* meta_state;
  const int32_t *cu, *slot, *init_mode;
  int N, H, Hk, Hv, Hs, write_state;
  int seg_len, n_segs, pass;   // sequence segments (0: whole sequences); pass 1 = segment summaries, 3 = outputs
  float* ws;                   // [N*Hs][n_segs][GP_WS_ELEMS] fp32: local end state + total decay per segment
  float scale;
  int64_t q_st, q_sh, k_st, k_sh, v_st, v_sh, g_st, g_sh, o_st, o_sh, st_ss, st_sh, meta_sh;
  const __nv_bfloat16* gnorm_w;   // per-head RMSNorm weight [128] of the 7B g_norm epilogue, or null
  float gnorm_eps;
  uint32_t relu_lim;              // bf16x2 lower limit applied to q and k: 0 = relu (7B feature map on raw
                                  // projection outputs), 0xff80ff80 (-inf) = identity
  float* sumsq;                   // f1: null, or [T, H] fp32 -- sum of squares of each bf16 output row (the branch
                                  // merge's RMS-norm partials), emitted by the O epilogue
  // balanced single-pass form (all null otherwise): plan header, per-CTA piece ranges and step
  // totals, piece table, hand-off flags of the fp32 buffers in ws (one per cut)
  int32_t* plan;
  const int32_t *cta_off, *cta_steps;
  const GlaPiece* pieces;
  int32_t* flags;
  const uint16_t* gate_table;     // null, or [65536] bf16: g = table[raw bf16 gate bits] -- the 7B gate
                                  // bf16(bf16(logsigmoid(gk)) / normalizer) is a pure function of a 16-bit
                                  // input, tabulated once per normalizer by hymeta_gla_gate_7b_g itself
                                  // (bit-identical by construction; the 128 KB table lives in L1 / L2)
};

__device__ __forceinline__ uint32_t hmul2_bf16(uint32_t a, uint32_t b) {   // exact products, one rounding each
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint32_t max_bf16x2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               :: "r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]),
                  "r"(r[6]), "r"(r[7]) : "memory");
}

#ifdef HYMETA_TRACE
__device__ long long* g_gla_trace = nullptr;   // [steps][16] clock64 stamps of CTA (0,0); tools/trace_gla_prefill.py
#define GP_TRACE(step, ev) do { if (g_gla_trace && blockIdx.x == 0 && blockIdx.y == 0 && (step) < 4096) \
    g_gla_trace[(step) * 16 + (ev)] = clock64(); } while (0)
#else
#define GP_TRACE(step, ev) do {} while (0)
#endif

template <bool GF32, bool GN>
__global__ void __launch_bounds__(GP_THREADS, 1)
gla_prefill_kernel(const GlaPrefillParams p, const __grid_constant__ CUtensorMap tmG,
                   const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler
  const int G = p.H / p.Hs;
  const bool summary = p.pass == 1;
  // ---- this CTA's work list -------------------------------------------------------------------
  GlaPiece own;                    // the block index's piece (grid launches) / first piece of the list
  own.pad = 0;
  int pi0 = 0, n_pieces = 1, cta_steps = 0;
  const bool table = p.plan != nullptr;   // balanced launch: `ctas` CTAs, each with a list from the plan kernel
  if (table) {
    // logical CTA index = arrival order, so the CTA a hand-off waits for has always started
    int* slot_l = reinterpret_cast<int*>(smem + GP_BAR + 208);
    if (tid == 0) *slot_l = atomicAdd(p.plan + 2, 1);
    __syncthreads();
    const int logical = *slot_l;
    if (logical >= p.plan[1]) return;
    pi0 = p.cta_off[logical];
    n_pieces = p.cta_off[logical + 1] - pi0;
    cta_steps = p.cta_steps[logical];
    if (n_pieces == 0) return;
    own = p.pieces[pi0];
  } else {
    const int hs = blockIdx.x, n = blockIdx.y;
    const int seq_begin = p.cu[n], seq_len = p.cu[n + 1] - seq_begin;
    const int slot = p.slot[n], mode = p.init_mode[n];
    // segment [seg0, seg0 + len) of the sequence (the whole sequence when seg_len == 0)
    const int seg = blockIdx.z, seg0 = seg * p.seg_len;
    if (seg > 0 && seg0 >= seq_len) return;
    const bool last_seg = p.seg_len == 0 || seg0 + p.seg_len >= seq_len;
    if (summary && last_seg) return;               // nothing follows: no summary needed
    const int ws_idx = (n * p.Hs + hs) * p.n_segs + seg;
    own.tok_begin = seq_begin + seg0;
    own.len = last_seg ? seq_len - seg0 : p.seg_len;
    own.hs = hs;
    own.in_kind = 0; own.in_idx = 0;
    if (p.pass == 3 && seg > 0) { own.in_kind = 3; own.in_idx = ws_idx - 1; }
    else if (!summary && seg == 0) {
      if (mode == 1 && slot >= 0) { own.in_kind = 1; own.in_idx = slot; }
      else if (mode == 2 && p.meta_state) own.in_kind = 2;
    }
    own.out_kind = 0; own.out_idx = 0;
    if (summary) { own.out_kind = 2; own.out_idx = ws_idx; }
    else if (last_seg && p.write_state && slot >= 0) { own.out_kind = 1; own.out_idx = slot; }
    cta_steps = (own.len + GC - 1) / GC;
    if (cta_steps == 0) {   // empty sequence: the slot receives the initial state unchanged
      if (own.out_kind == 1 && warp >= 4 && warp < 8) {
        const __nv_bfloat16* init_src = nullptr;
        if (own.in_kind == 1) init_src = p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh;
        else if (own.in_kind == 2) init_src = p.meta_state + (int64_t)hs * p.meta_sh;
        __nv_bfloat16* final_dst = p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh;
        if (final_dst != init_src) {
          const int d = tid - 128;
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            uint4 t = make_uint4(0, 0, 0, 0);
            if (init_src) t = ldg_v4(init_src + d * kD + 8 * c);
            stg_v4(final_dst + d * kD + 8 * c, t);
          }
        }
      }
      return;
    }
  }
  auto piece = [&](int k) -> GlaPiece {
    if (!table || k == 0) return own;
    const int4* src = reinterpret_cast<const int4*>(p.pieces + pi0 + k);
    const int4 a = __ldg(src), b = __ldg(src + 1);
    GlaPiece r;
    r.tok_begin = a.x; r.len = a.y; r.hs = a.z; r.in_kind = a.w;
    r.in_idx = b.x; r.out_kind = b.y; r.out_idx = b.z; r.pad = 0;
    return r;
  };

  const uint32_t sbase = smem_u32(smem);
  const uint32_t sQ = sbase + GP_SQ, sS = sbase + GP_SS, sT = sbase + GP_ST;
  const uint32_t bars = sbase + GP_BAR;
  const uint32_t bar_kf = bars, bar_ke = bar_kf + 8 * GP_KS, bar_qf = bar_ke + 8 * GP_KS,
                 bar_a = bar_qf + 16, bar_p = bar_a + 16, bar_o = bar_p + 16, bar_x = bar_o + 16,
                 bar_xf = bar_x + 16, bar_raw = bar_xf + 16, bar_fin = bar_raw + 8 * GP_RS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + GP_BAR + 200);

  auto init_bars = [&](bool again) {
    if (again)   // a piece is over: every phase has completed, no asynchronous arrival is pending
      for (uint32_t b = bar_kf; b <= bar_fin; b += 8) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" :: "r"(b) : "memory");
    for (int s = 0; s < GP_KS; ++s) { mbar_init(bar_kf + 8 * s, 64); mbar_init(bar_ke + 8 * s, summary ? 128 : 257); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_qf + 8 * i, 128); mbar_init(bar_a + 8 * i, 1);
      mbar_init(bar_p + 8 * i, 128);  mbar_init(bar_o + 8 * i, 1);
    }
    mbar_init(bar_x, 128); mbar_init(bar_x + 8, 128); mbar_init(bar_xf, 1); mbar_init(bar_xf + 8, 1);
    for (int s = 0; s < GP_RS; ++s) mbar_init(bar_raw + 8 * s, 1);
    mbar_init(bar_fin, 1);
    fence_mbar_init();
  };
  // Between the pieces of a work list the pipeline drains and restarts from step 0 (a few steps of
  // ramp per piece, ~1 % at config 2): every role's loops stay those of a single chain.
  auto piece_end = [&]() {
    tc_fence_before();
    asm volatile("bar.sync 0;" ::: "memory");     // every role is done with the piece
    if (tid == 0) init_bars(true);
    asm volatile("bar.sync 0;" ::: "memory");
    tc_fence_after();
  };
  if (tid == 0) {
    if (sbase & 1023) asm volatile("trap;");
    init_bars(false);
  }
  if (GN && tid >= 64 && tid < 80)   // g_norm weight -> smem once (read as broadcast words by every row owner)
    *reinterpret_cast<uint4*>(smem + GP_GW + (tid - 64) * 16) = ldg_nc_v4(p.gnorm_w + (tid - 64) * 8);
  if (warp == 2) tmem_alloc<512>(smem_u32(tmem_slot));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // TMEM columns: O0 [0,128)  O1 [128,256)  A0 [256,272)  A1 [272,288); P_i aliases A_i
  const uint32_t tO = tmem, tA = tmem + 256;

  // 512 threads x 128 registers at launch; the state warpgroup takes what the others give up
  if (warp < 4) {
   setmaxnreg_dec<80>();
   if (warp < 2) {
    // ================================================================== gates (thread = 2 channels)
    // Straight-line code is kept small on purpose: five roles share the SM's instruction cache.
    const int c = 2 * tid;
    constexpr int GB = GF32 ? 4 : 2;
    const uint32_t kbase = (uint32_t)((c >> 6) * 2048 + (c & 7) * 2), kcx = (uint32_t)((c >> 3) & 7);
    const float lscale = __log2f(p.scale);
    // raw g / k of the next GP_RS steps are TMA-loaded into a small ring ([16][128] rows, no swizzle)
    // and v straight into its SWIZZLE_128B stage tile: no load instructions in this role
    constexpr uint32_t RAW_BYTES = GC * kD * (GB + 2);
    const uint8_t* rawp = smem + GP_RAW + c * GB;
    const uint8_t* rawkp = smem + GP_RAW + GP_RAW_K + c * 2;
    if (tid == 0) { tma_prefetch_desc(&tmG); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); }
    for (int pi = 0; pi < n_pieces; ++pi) {
    const GlaPiece pc = piece(pi);
    const int n_steps = (pc.len + GC - 1) / GC, len = pc.len, tok_begin = pc.tok_begin;
    const int hk = pc.hs / (p.Hs / p.Hk), hv = pc.hs / (p.Hs / p.Hv);
    auto fetch = [&](int i) {
      if (i < n_steps && tid == 0) {
        const uint32_t bar = bar_raw + 8 * (i % GP_RS), dst = sbase + GP_RAW + (i % GP_RS) * GP_RAW_STAGE;
        mbar_expect_tx(bar, RAW_BYTES);
        tma_load_3d(dst, &tmG, bar, 0, hk, tok_begin + i * GC);
        tma_load_3d(dst + GP_RAW_K, &tmK, bar, 0, hk, tok_begin + i * GC);
      }
    };
#pragma unroll
    for (int i = 0; i < GP_RS; ++i) fetch(i);
    for (int i = 0; i < n_steps; ++i) {
      const int st = i % GP_KS, clen = len - i * GC;
      const uint32_t so = (uint32_t)((i % GP_RS) * GP_RAW_STAGE);
      uint8_t* stage = smem + GP_ST + st * GP_STAGE;
      if (tid == 0) GP_TRACE(i, 15);
      mbar_wait(bar_raw + 8 * (i % GP_RS), (i / GP_RS) & 1);
      if (tid == 0) GP_TRACE(i, 14);
      // cumulative gates in the log2 domain; rows past the sequence end (t >= clen) hold another
      // sequence's tokens: they neither decay nor contribute (k~ = 0)
      float2 b[GC];
      uint32_t kk[GC];
      // every load is issued before the first use (no branch between them: the ragged tail is
      // handled by selects, so the 32 shared-memory reads overlap instead of serialising)
#pragma unroll
      for (int t = 0; t < GC; ++t) {
        if (GF32) b[t] = *reinterpret_cast<const float2*>(rawp + so + t * 512);
        else {
          const uint32_t u = *reinterpret_cast<const uint32_t*>(rawp + so + t * 256);
          b[t] = make_float2(__uint_as_float(u & 0xffffu), __uint_as_float(u >> 16));   // raw bits for now
        }
        kk[t] = *reinterpret_cast<const uint32_t*>(rawkp + so + t * 256);
      }
      if (!GF32) {
        if (p.gate_table) {   // 32 independent 2-byte gathers per thread and step, hot entries in L1
#pragma unroll
          for (int t = 0; t < GC; ++t) {
            const uint32_t lo = __ldg(p.gate_table + __float_as_uint(b[t].x));
            const uint32_t hi = __ldg(p.gate_table + __float_as_uint(b[t].y));
            b[t] = make_float2(__uint_as_float(lo << 16), __uint_as_float(hi << 16));
          }
        } else {
#pragma unroll
          for (int t = 0; t < GC; ++t)
            b[t] = make_float2(__uint_as_float(__float_as_uint(b[t].x) << 16), __uint_as_float(__float_as_uint(b[t].y) << 16));
        }
      }
      float bx = 0.f, by = 0.f;
#pragma unroll
      for (int t = 0; t < GC; ++t) {
        const bool in = t < clen;
        bx = fmaf(in ? b[t].x : 0.f, 1.4426950408889634f, bx);
        by = fmaf(in ? b[t].y : 0.f, 1.4426950408889634f, by);
        kk[t] = in ? max_bf16x2(kk[t], p.relu_lim) : 0u;
        b[t] = make_float2(bx, by);
      }
      const float2 r = b[GC / 2 - 1];
      const float rfx = r.x - lscale, rfy = r.y - lscale;
      if (tid == 0) GP_TRACE(i, 2);
      if (i >= GP_KS) mbar_wait(bar_ke + 8 * st, ((i / GP_KS) - 1) & 1);
      if (clen >= GC) {
        if (tid == 0) {
          const uint32_t dst = sT + st * GP_STAGE + GP_ST_V;
          mbar_expect_tx_only(bar_kf + 8 * st, GC * kD * 2);
          tma_load_3d(dst, &tmV, bar_kf + 8 * st, 0, hv, tok_begin + i * GC);
          tma_load_3d(dst + 2048, &tmV, bar_kf + 8 * st, 64, hv, tok_begin + i * GC);
        }
      } else {   // ragged last step: copy by hand, zero rows past the sequence end
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int id = tid + 64 * j, t = id >> 4, vc = id & 15;
          uint4 x = make_uint4(0, 0, 0, 0);
          if (t < clen) x = ldg_nc_v4(p.v + (int64_t)(tok_begin + i * GC + t) * p.v_st + (int64_t)hv * p.v_sh + vc * 8);
          *reinterpret_cast<uint4*>(stage + GP_ST_V + (vc >> 3) * 2048 + sw128_chunk(t, vc & 7)) = x;
        }
      }
      uint8_t* sF = stage + GP_ST_F + c * 4;
      uint8_t* sK = stage + GP_ST_K + kbase;
#pragma unroll
      for (int t = 0; t < GC; ++t) {
        if (GP_EXP & 8) {
          *reinterpret_cast<float2*>(sF + t * (kD * 4)) = make_float2(b[t].x, b[t].y);
          *reinterpret_cast<uint32_t*>(sK + t * 128 + ((kcx ^ (t & 7)) << 4)) = kk[t];
          continue;
        }
        *reinterpret_cast<float2*>(sF + t * (kD * 4)) = make_float2(exp2f(b[t].x - rfx), exp2f(b[t].y - rfy));
        *reinterpret_cast<uint32_t*>(sK + t * 128 + ((kcx ^ (t & 7)) << 4)) =
            pack_bf16(bf16_lo(kk[t]) * exp2f(r.x - b[t].x), bf16_hi(kk[t]) * exp2f(r.y - b[t].y));
      }
      float* sD = reinterpret_cast<float*>(stage + GP_ST_D);
      *reinterpret_cast<float2*>(sD + c) = make_float2(exp2f(b[GC - 1].x - r.x), exp2f(b[GC - 1].y - r.y));
      *reinterpret_cast<float2*>(sD + kD + c) = make_float2(exp2f(r.x), exp2f(r.y));
      fence_proxy_async_smem();
      mbar_arrive(bar_kf + 8 * st);
      // The raw slot is refilled only now: the stores above consumed every value loaded from it, so
      // no shared-memory read of this step can still be in flight when the TMA write lands (a
      // barrier right after the loads is NOT enough -- the compiler schedules their first use
      // behind it, and under LSU pressure the last rows were read after the refill).
      asm volatile("bar.sync 1, 64;" ::: "memory");   // both gate warps are done with the raw slot
      fetch(i + GP_RS);
      if (tid == 0) GP_TRACE(i, 0);
    }
    piece_end();
    }
   } else if (warp == 3) {
    // idle: g / k / v arrive by TMA
    for (int pi = 0; pi < n_pieces; ++pi) piece_end();
   } else {
    // ================================================================== MMA issuer
    for (int pi = 0; pi < n_pieces; ++pi) {
    const int n_steps = (piece(pi).len + GC - 1) / GC;
    if (!summary && n_steps > 0 && elect_one_sync()) {
      const uint32_t idesc_a = umma_idesc_bf16(128, 16, 0, 0);
      const uint32_t idesc_o = umma_idesc_bf16(128, 128, 0, 1);
      // descriptors differ from these templates only in the 16-byte-granular start address
      const uint64_t dk = umma_desc_sw128(0, 16, 1024);          // K-major tiles (q~, k~ as B of A)
      const uint64_t dmn = umma_desc_sw128(0, 2048, 1024);       // MN-major 16-row tiles (k~ as A of Y, v)
      const uint64_t dx = umma_desc_sw128(0, 16384, 1024);       // MN-major X (two 64-wide halves)
      auto issue_a = [&](int i) {                   // A = q~ k~^T
        const int w = i & 1;
        const uint32_t qb = (sQ + w * 32768) >> 4, kb = (sT + (i % GP_KS) * GP_STAGE + GP_ST_K) >> 4;
#pragma unroll 1
        for (int ks = 0; ks < 8; ++ks)
          umma_bf16(tA + 16 * w, dk + qb + (ks >> 2) * 1024 + (ks & 3) * 2, dk + kb + (ks >> 2) * 128 + (ks & 3) * 2,
                    idesc_a, ks > 0);
        umma_commit(bar_a + 8 * w);
      };
      mbar_wait(bar_kf, 0);
      mbar_wait(bar_qf, 0);
      tc_fence_after();
      issue_a(0);
      for (int i = 0; i < n_steps; ++i) {
        const int w = i & 1, st = i % GP_KS;
        const uint32_t vb = (sT + st * GP_STAGE + GP_ST_V) >> 4;
        GP_TRACE(i, 9);
        mbar_wait(bar_x + 8 * w, (i >> 1) & 1);
        tc_fence_after();
        GP_TRACE(i, 10);
        {                                                // O = q~ X
          const uint32_t qb = (sQ + w * 32768) >> 4, xb = (sS + w * 32768) >> 4;
#pragma unroll 1
          for (int ks = 0; ks < 8; ++ks)
            umma_bf16(tO + 128 * w, dk + qb + (ks >> 2) * 1024 + (ks & 3) * 2, dx + xb + ks * 128, idesc_o, ks > 0);
          umma_commit(bar_xf + 8 * w);                   // X buffer w free once these complete
        }
        // A(i+1) goes ahead of P v (i) only when its operands are already there; otherwise P v (i)
        // and the epilogue behind it are not held up by the next tile's builder
        bool a_done = i + 1 >= n_steps;
        if (!a_done && mbar_test(bar_qf + 8 * (w ^ 1), ((i + 1) >> 1) & 1) &&
            mbar_test(bar_kf + 8 * ((i + 1) % GP_KS), ((i + 1) / GP_KS) & 1)) {
          tc_fence_after();
          issue_a(i + 1);
          a_done = true;
        }
        GP_TRACE(i, 11);
        mbar_wait(bar_p + 8 * w, (i >> 1) & 1);
        tc_fence_after();
        GP_TRACE(i, 12);
        umma_bf16_ts(tO + 128 * w, tA + 16 * w, dmn + vb, idesc_o, 1);   // O += P v
        umma_commit(bar_o + 8 * w);
        umma_commit(bar_ke + 8 * st);                    // k~, v of step i consumed
        if (!a_done) {
          mbar_wait(bar_kf + 8 * ((i + 1) % GP_KS), ((i + 1) / GP_KS) & 1);
          mbar_wait(bar_qf + 8 * (w ^ 1), ((i + 1) >> 1) & 1);
          tc_fence_after();
          issue_a(i + 1);
        }
        GP_TRACE(i, 13);
      }
      // every MMA of the piece has completed (and with it every commit it was to signal) before the
      // barriers are torn down
      umma_commit(bar_fin);
      mbar_wait(bar_fin, 0);
    }
    __syncwarp();
    piece_end();
    }
   }
  } else if (warp < 8) {
    setmaxnreg_inc<208>();
    // ================================================================== state (fp32 master in registers)
    // TMEM -> register traffic is 64 B/clk per SM and the O epilogue already needs most of it, so the
    // state never enters TMEM: Y = X + k~^T v runs on mma.sync with the accumulator fragments as the
    // master (warp = 32 d_k rows x 128 d_v), is rescaled in registers and its bf16 copy goes to smem.
    // The chain needs only the gate ring and a free X buffer, so it runs ahead of the MMA issuer.
    const int ws = warp - 4, gid = lane >> 2, tig = lane & 3;
    const int row0 = 32 * ws + gid;                 // rows row0 + 8j, j = 0..3 (m-tile j>>1, half j&1)
    // ldmatrix row addresses inside a stage (k~ / v tiles: [2 halves][16 tokens][64] SWIZZLE_128B)
    const int lt = (lane & 7) + 8 * ((lane >> 4) & 1);          // A: token row, d_k block (lane>>3)&1
    const uint32_t a_row = (uint32_t)(lt * 128), a_x = (uint32_t)(lt & 7);
    const int a_blk = 4 * ws + ((lane >> 3) & 1);               // 8-wide d_k block of m-tile 0 (+2 for m-tile 1)
    const int bt = (lane & 7) + 8 * ((lane >> 3) & 1);          // B: token row, d_v block pair lane>>4
    const uint32_t b_row = (uint32_t)(bt * 128), b_x = (uint32_t)(bt & 7);
    const float* sD0 = reinterpret_cast<const float*>(smem + GP_ST + GP_ST_D);
    float S[2][16][4];
    for (int pi = 0; pi < n_pieces; ++pi) {
    const GlaPiece pc = piece(pi);
    const int n_steps = (pc.len + GC - 1) / GC;
    // ---- start state of the piece
    const __nv_bfloat16* init_src = nullptr;
    const float* ws_in = nullptr;
    if (pc.in_kind == 1) init_src = p.state + (int64_t)pc.in_idx * p.st_ss + (int64_t)pc.hs * p.st_sh;
    else if (pc.in_kind == 2) init_src = p.meta_state + (int64_t)pc.hs * p.meta_sh;
    else if (pc.in_kind == 3) {
      ws_in = p.ws + (int64_t)pc.in_idx * GP_WS_ELEMS;
      if (p.flags) {   // handed over by the CTA that owns the head of this chain (it ran it first)
        if (tid == 128) {
          int v;
          do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p.flags + pc.in_idx) : "memory");
          } while (v == 0);
        }
        asm volatile("bar.sync 2, 128;" ::: "memory");
      }
    }
    __nv_bfloat16* final_dst = pc.out_kind == 1
        ? p.state + (int64_t)pc.out_idx * p.st_ss + (int64_t)pc.hs * p.st_sh : nullptr;
    float* ws_out = pc.out_kind == 2 ? p.ws + (int64_t)pc.out_idx * GP_WS_ELEMS : nullptr;
    float Ptot[4], Din[4];   // total decay of the segment (summary pass); decay owed by a handed-over state
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int nt = 0; nt < 16; ++nt)
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          const int off = (row0 + 16 * mt + 8 * hf) * kD + 8 * nt + 2 * tig;
          if (ws_in) {
            const float2 f = __ldcg(reinterpret_cast<const float2*>(ws_in + off));
            S[mt][nt][2 * hf] = f.x; S[mt][nt][2 * hf + 1] = f.y;
          } else {
            uint32_t u = 0;
            if (init_src) u = *reinterpret_cast<const uint32_t*>(init_src + off);
            S[mt][nt][2 * hf] = bf16_lo(u); S[mt][nt][2 * hf + 1] = bf16_hi(u);
          }
        }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      Ptot[j] = 1.f;
      // at a cut the state travels WITHOUT its last step's decay, which is folded into the first
      // rescale here exactly as the uncut chain folds it: S * (d_i * e_{i+1}), bit for bit
      Din[j] = (ws_in && p.flags) ? __ldcg(ws_in + kD * kD + row0 + 8 * j) : 1.f;
    }
    for (int i = -1; i < n_steps; ++i) {
      const bool first = i < 0, last = i + 1 == n_steps;
      if (!first && !(GP_EXP & 16)) {
        const uint32_t stg = sT + (i % GP_KS) * GP_STAGE;      // kf(i) was awaited in pass i-1
        uint32_t af[2][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          const int blk = a_blk + 2 * mt;                       // 0..15
          ldmatrix_x4_trans(af[mt][0], af[mt][1], af[mt][2], af[mt][3],
                            stg + GP_ST_K + (blk >> 3) * 2048 + a_row + ((((uint32_t)blk & 7) ^ a_x) << 4));
        }
#pragma unroll
        for (int np = 0; np < 8; ++np) {                        // d_v blocks 2np, 2np+1
          const int blk = 2 * np + (lane >> 4);
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(b0, b1, b2, b3,
                            stg + GP_ST_V + (blk >> 3) * 2048 + b_row + ((((uint32_t)blk & 7) ^ b_x) << 4));
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) {
            mma_bf16_16816(S[mt][2 * np], af[mt], b0, b1);
            mma_bf16_16816(S[mt][2 * np + 1], af[mt], b2, b3);
          }
        }
      }
      float D[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) D[j] = first ? Din[j] : sD0[(i % GP_KS) * (GP_STAGE / 4) + row0 + 8 * j];
      if (!last) {
        const int sn = (i + 1) % GP_KS;
        mbar_wait(bar_kf + 8 * sn, ((i + 1) / GP_KS) & 1);
#pragma unroll
        for (int j = 0; j < 4; ++j) D[j] *= sD0[sn * (GP_STAGE / 4) + kD + row0 + 8 * j];
      }
      if (!first) mbar_arrive(bar_ke + 8 * (i % GP_KS));        // k~, v, D of step i consumed
      const bool handoff = last && ws_out && p.flags;           // cut: the decay travels beside the state
      if (!handoff) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < 16; ++nt) {
            S[mt][nt][0] *= D[2 * mt]; S[mt][nt][1] *= D[2 * mt];
            S[mt][nt][2] *= D[2 * mt + 1]; S[mt][nt][3] *= D[2 * mt + 1];
          }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) Ptot[j] = handoff ? D[j] : Ptot[j] * D[j];
      if (summary) {
        if (last) {   // local end state (zero start) and total decay of this segment
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 16; ++nt)
#pragma unroll
              for (int hf = 0; hf < 2; ++hf)
                *reinterpret_cast<float2*>(ws_out + (row0 + 16 * mt + 8 * hf) * kD + 8 * nt + 2 * tig) =
                    make_float2(S[mt][nt][2 * hf], S[mt][nt][2 * hf + 1]);
          if (tig == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) ws_out[kD * kD + row0 + 8 * j] = Ptot[j];
          }
        }
      } else if (!last) {
        // X_{i+1} -> buffer (i+1)&1, last read by q~ X of step i-1
        if (i >= 1) mbar_wait(bar_xf + 8 * ((i + 1) & 1), ((i - 1) >> 1) & 1);
        // rows row0 + 8j all have (row & 7) == gid: eight slot bases, everything else is immediate
        uint8_t* sdst = smem + GP_SS + ((i + 1) & 1) * 32768 + row0 * 128 + tig * 4;
#pragma unroll
        for (int k = 0; k < ((GP_EXP & 32) ? 0 : 8); ++k) {
          uint8_t* bk = sdst + (((uint32_t)(k ^ gid) & 7) << 4);
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int hf = 0; hf < 2; ++hf)
#pragma unroll
              for (int h2 = 0; h2 < 2; ++h2)
                *reinterpret_cast<uint32_t*>(bk + (16 * mt + 8 * hf) * 128 + h2 * 16384) =
                    pack_bf16(S[mt][8 * h2 + k][2 * hf], S[mt][8 * h2 + k][2 * hf + 1]);
        }
        fence_proxy_async_smem();
        mbar_arrive(bar_x + 8 * ((i + 1) & 1));   // one barrier per buffer: the state may run a full buffer ahead
        if (tid == 128 && i >= 0) GP_TRACE(i, 3);
      } else if (ws_out) {   // cut: fp32 state (without the last decay) and that decay, then the flag
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < 16; ++nt)
#pragma unroll
            for (int hf = 0; hf < 2; ++hf)
              __stcg(reinterpret_cast<float2*>(ws_out + (row0 + 16 * mt + 8 * hf) * kD + 8 * nt + 2 * tig),
                     make_float2(S[mt][nt][2 * hf], S[mt][nt][2 * hf + 1]));
        if (tig == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) __stcg(ws_out + kD * kD + row0 + 8 * j, Ptot[j]);
        }
        __threadfence();
        asm volatile("bar.sync 2, 128;" ::: "memory");
        if (tid == 128) {
          const int one = 1;
          asm volatile("st.release.gpu.global.s32 [%0], %1;" :: "l"(p.flags + pc.out_idx), "r"(one) : "memory");
        }
      } else if (final_dst) {
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
          for (int nt = 0; nt < 16; ++nt)
#pragma unroll
            for (int hf = 0; hf < 2; ++hf)
              *reinterpret_cast<uint32_t*>(final_dst + (row0 + 16 * mt + 8 * hf) * kD + 8 * nt + 2 * tig) =
                  pack_bf16(S[mt][nt][2 * hf], S[mt][nt][2 * hf + 1]);
      }
    }
    piece_end();
    }
  } else {
    setmaxnreg_dec<112>();
    // ================================================================== q tiles (even / odd steps)
    const int w = (warp - 8) >> 2;
    const int tq = tid - 256 - w * 128;             // 0..127
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const int qc = tq & 15, qt0 = tq >> 4;          // builder view: 16-byte chunk qc, tokens qt0, qt0+8
    const int r_t = tq & 15;                        // row-owner view: token (head slot = tq >> 4)
    const uint32_t tAw = tA + 16 * w + lane_off, tOw = tO + 128 * w + lane_off;
    // q~ row (hh*16 + t) of 16-byte chunk qc: byte offsets of tokens qt0 / qt0+8 for head slot 0
    uint8_t* sQw = smem + GP_SQ + w * 32768 + (qc >> 3) * 16384;
    const uint32_t qo0 = (uint32_t)(qt0 * 128 + (((qc ^ qt0) & 7) << 4)), qo1 = qo0 + 1024;   // (qt0+8)&7 == qt0&7
    uint4 qv[16];
    for (int pi = 0; pi < n_pieces; ++pi) {
    const GlaPiece pc = piece(pi);
    const int n_steps = (pc.len + GC - 1) / GC, len = pc.len, tok_begin = pc.tok_begin, h0 = pc.hs * G;
    const __nv_bfloat16* qp = p.q + (int64_t)(tok_begin + qt0) * p.q_st + (int64_t)h0 * p.q_sh + qc * 8;
    const int q_steps = summary ? 0 : n_steps;      // the summary pass has no queries
    int cnt = 0;
    for (int i = w - 2; i < q_steps; i += 2) {
      const int clen = len - i * GC;
      if (i >= 0) {
        const int st = i % GP_KS;
        // ---- build q~ from the registers prefetched one tile ahead (buffer w is free: this
        //      warpgroup saw bar_o of step i-2 in its epilogue)
        if (tq == 0) GP_TRACE(i, 4);
        mbar_wait(bar_kf + 8 * st, (i / GP_KS) & 1);
        const float* sF = reinterpret_cast<const float*>(smem + GP_ST + st * GP_STAGE + GP_ST_F) + qt0 * kD + qc * 8;
        float e0[8], e1[8];
        *reinterpret_cast<float4*>(&e0[0]) = *reinterpret_cast<const float4*>(sF);
        *reinterpret_cast<float4*>(&e0[4]) = *reinterpret_cast<const float4*>(sF + 4);
        *reinterpret_cast<float4*>(&e1[0]) = *reinterpret_cast<const float4*>(sF + 8 * kD);
        *reinterpret_cast<float4*>(&e1[4]) = *reinterpret_cast<const float4*>(sF + 8 * kD + 4);
        mbar_arrive(bar_ke + 8 * st);               // factor table consumed
#pragma unroll
        for (int it = 0; it < 16; ++it) {
          const float* e = (it & 1) ? e1 : e0;
          const uint4 v4 = qv[it];
          uint4 o4;
          // relu(q) * e == max(q * e, 0) because e > 0: one packed max per word (identity for -inf)
          o4.x = max_bf16x2(pack_bf16(bf16_lo(v4.x) * e[0], bf16_hi(v4.x) * e[1]), p.relu_lim);
          o4.y = max_bf16x2(pack_bf16(bf16_lo(v4.y) * e[2], bf16_hi(v4.y) * e[3]), p.relu_lim);
          o4.z = max_bf16x2(pack_bf16(bf16_lo(v4.z) * e[4], bf16_hi(v4.z) * e[5]), p.relu_lim);
          o4.w = max_bf16x2(pack_bf16(bf16_lo(v4.w) * e[6], bf16_hi(v4.w) * e[7]), p.relu_lim);
          if (!(GP_EXP & 4)) *reinterpret_cast<uint4*>(sQw + (it >> 1) * 2048 + ((it & 1) ? qo1 : qo0)) = o4;
        }
        fence_proxy_async_smem();
        tc_fence_before();                          // orders the previous epilogue's TMEM reads too
        mbar_arrive(bar_qf + 8 * w);
        if (tq == 0) GP_TRACE(i, 5);
      }
      if (i + 2 < q_steps) {                        // next tile: in flight under the P conversion and epilogue
        const int rem = len - (i + 2) * GC - qt0;   // token qt0 + 8e valid iff 8e < rem
        const __nv_bfloat16* src = qp + (int64_t)(i + 2) * GC * p.q_st;
#pragma unroll
        for (int it = 0; it < 16; ++it) {
          qv[it] = make_uint4(0, 0, 0, 0);
          if ((it >> 1) < G && (it & 1) * 8 < rem)
            qv[it] = ldg_nc_v4(src + (int64_t)(it >> 1) * p.q_sh + (int64_t)((it & 1) * 8) * p.q_st);
        }
      }
      if (i >= 0) {
        // ---- causal A -> bf16 P, in place in TMEM (thread = row)
        mbar_wait(bar_a + 8 * w, cnt & 1);
        tc_fence_after();
        if (tq == 0) GP_TRACE(i, 6);
        {
          float a[16];
          tmem_ld16(tAw, a);
          uint32_t pk[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            pk[j] = pack_bf16((2 * j <= r_t) ? a[2 * j] : 0.f, (2 * j + 1 <= r_t) ? a[2 * j + 1] : 0.f);
          tmem_st8(tAw, pk);
          tmem_wait_st();
          tc_fence_before();
          mbar_arrive(bar_p + 8 * w);
          if (tq == 0) GP_TRACE(i, 7);
        }
        // ---- epilogue: O rows -> global, 256 contiguous bytes per (token, head)
        mbar_wait(bar_o + 8 * w, cnt & 1);
        tc_fence_after();
        if (tq == 0) GP_TRACE(i, 8);
        {
          // 7B g_norm epilogue (model_for_7B/gla_attention.py:177,210): vLLM RMSNorm over the row's
          // 128 outputs of the bf16-rounded o, y = bf16(bf16(o * rstd) * w).  The thread owns the
          // whole row, so this is a first TMEM pass for the sum of squares and no exchange.
          // x = bf16(o) is what both the sum of squares and the scaling use: pass 1 rounds each
          // 32-column chunk once, packs it (bf16x2) back over the consumed columns of the O region
          // and accumulates the squares; pass 2 reads the 64 packed columns, scales in fp32, rounds,
          // and applies the weight with one packed bf16 multiply (exact product, one rounding --
          // the same value as the fp32 multiply + round of the reference op).
          float rstd = 0.f;
          if (GN) {
            float ss = 0.f;
_Pragma(GP_STR(unroll GP_GN_UNROLL))
            for (int cc = 0; cc < 8; ++cc) {   // 16 columns at a time: the q prefetch holds 64 registers here
              float t[16];
              tmem_ld16(tOw + 16 * cc, t);
              uint32_t xb[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                xb[j] = pack_bf16(t[2 * j], t[2 * j + 1]);
                ss = fmaf(bf16_lo(xb[j]), bf16_lo(xb[j]), ss);
                ss = fmaf(bf16_hi(xb[j]), bf16_hi(xb[j]), ss);
              }
              tmem_st8(tOw + 8 * cc, xb);
            }
            tmem_wait_st();
            rstd = rsqrtf(ss * (1.f / kD) + p.gnorm_eps);
          }
          const uint32_t* gw = reinterpret_cast<const uint32_t*>(smem + GP_GW);
          float ssq = 0.f;
#pragma unroll 1
          for (int cc = 0; cc < ((GP_EXP & 1) ? 0 : 4); ++cc) {
            uint32_t ob[16];   // 32 output columns, packed bf16x2
            if (GN) {
              float t[16];
              tmem_ld16(tOw + 16 * cc, t);
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t x = __float_as_uint(t[j]);
                ob[j] = hmul2_bf16(pack_bf16(bf16_lo(x) * rstd, bf16_hi(x) * rstd), gw[16 * cc + j]);
              }
            } else {
#pragma unroll
              for (int hf = 0; hf < 2; ++hf) {
                float t[16];
                tmem_ld16(tOw + 32 * cc + 16 * hf, t);
#pragma unroll
                for (int j = 0; j < 8; ++j) ob[8 * hf + j] = pack_bf16(t[2 * j], t[2 * j + 1]);
              }
            }
            if (p.sumsq) {
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                ssq = fmaf(bf16_lo(ob[j]), bf16_lo(ob[j]), ssq);
                ssq = fmaf(bf16_hi(ob[j]), bf16_hi(ob[j]), ssq);
              }
            }
            // 32 rows x 64 bytes staged per warp (chunks XOR-swizzled), read back so that four lanes
            // cover 64 contiguous bytes of one (token, head) row: 8 half-lines per store instruction
            // instead of 32 scattered 16-byte pieces, which kept the LSU -- and every other role's
            // shared-memory and mbarrier traffic behind it -- busy for ~2300 cycles per step
            uint8_t* stg = smem + GP_OST + (warp - 8) * 2048;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              *reinterpret_cast<uint4*>(stg + lane * 64 + (((j ^ (lane >> 1)) & 3) << 4)) =
                  make_uint4(ob[4 * j], ob[4 * j + 1], ob[4 * j + 2], ob[4 * j + 3]);
            __syncwarp();
            if (!(GP_EXP & 2)) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int row = 8 * j + (lane >> 2), c = lane & 3;
                const int rq = (warp & 3) * 32 + row, hh2 = rq >> 4, t2 = rq & 15;
                const uint4 o4 = *reinterpret_cast<const uint4*>(stg + row * 64 + (((c ^ (row >> 1)) & 3) << 4));
                if (hh2 < G && t2 < clen)
                  stg_v4(p.o + (int64_t)(tok_begin + i * GC + t2) * p.o_st + (int64_t)(h0 + hh2) * p.o_sh + 32 * cc + 8 * c, o4);
              }
            }
            __syncwarp();
          }
          if (p.sumsq && (tq >> 4) < G && r_t < clen)
            p.sumsq[(int64_t)(tok_begin + i * GC + r_t) * p.H + h0 + (tq >> 4)] = ssq;
        }
        ++cnt;
      }
    }
    piece_end();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

// Pass 2 of the segmented prefill: per chain, turn the segment summaries (local end state L_p from a
// zero start, total decay P_p) into segment start states, in place:  S_{p+1} = P_p (.) S_p + L_p,
// S_0 = the chain's initial state.  Slot p then holds the start state of segment p+1.
__global__ void __launch_bounds__(256) gla_prefill_scan_kernel(const GlaPrefillParams p) {
  const int hs = blockIdx.x, n = blockIdx.y;
  const int seq_len = p.cu[n + 1] - p.cu[n];
  const int n_seg = (seq_len + p.seg_len - 1) / p.seg_len;
  if (n_seg <= 1) return;
  const int slot = p.slot[n], mode = p.init_mode[n];
  const __nv_bfloat16* init_src = nullptr;
  if (mode == 1 && slot >= 0) init_src = p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh;
  else if (mode == 2 && p.meta_state) init_src = p.meta_state + (int64_t)hs * p.meta_sh;
  float* ws = p.ws + (int64_t)(n * p.Hs + hs) * p.n_segs * GP_WS_ELEMS;
  const int e0 = (blockIdx.z * 256 + threadIdx.x) * 4;      // 4 consecutive d_v of one d_k row
  const int row = e0 >> 7;
  float4 S = make_float4(0.f, 0.f, 0.f, 0.f);
  if (init_src) {
    const uint2 u = *reinterpret_cast<const uint2*>(init_src + e0);
    S = make_float4(bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y));
  }
  for (int sg = 0; sg + 1 < n_seg; ++sg) {
    float* w = ws + (int64_t)sg * GP_WS_ELEMS;
    const float P = w[kD * kD + row];
    const float4 L = *reinterpret_cast<const float4*>(w + e0);
    S = make_float4(fmaf(P, S.x, L.x), fmaf(P, S.y, L.y), fmaf(P, S.z, L.z), fmaf(P, S.w, L.w));
    *reinterpret_cast<float4*>(w + e0) = S;
  }
}

// Work plan of the balanced single-pass form (one CTA, every launch).  The (sequence, state-head)
// chains are laid end to end on a line of 16-token steps and the line is cut into `ncta` equal
// ranges, one per SM, so chains > SMs no longer run as whole waves (256 chains on 148 SMs: 1024 ->
// 886 steps per SM).  A chain that straddles a cut becomes two pieces: its head A in CTA j, its tail
// B in CTA j+1, the fp32 state handed over through buffer j of the workspace.  Every CTA walks its
// range BACKWARDS -- A first, whole chains, B last: A has no dependency, and B starts no earlier than
// (range - |B|) >= |A| steps into its CTA, when A is long done.  That needs range >= longest chain;
// otherwise (a few very long chains among short ones, or no more chains than SMs) whole chains are
// dealt round-robin.  Either way the main kernel is launched with one persistent CTA per SM.
__global__ void __launch_bounds__(1024) gla_prefill_plan_kernel(
    const GlaPrefillParams p, int ncta, int32_t* cs, int32_t* cta_off, int32_t* cta_steps, GlaPiece* pieces) {
  __shared__ int s_sum[1024], s_max[1024], s_cnt[1024];
  const int tid = threadIdx.x, N = p.N, Hs = p.Hs;
  const int per = (N + 1023) / 1024, n0 = min(tid * per, N), n1 = min(n0 + per, N);
  int sum = 0, mx = 0;
  for (int n = n0; n < n1; ++n) {
    const int st = (p.cu[n + 1] - p.cu[n] + GC - 1) / GC;
    sum += st; mx = max(mx, st);
  }
  s_sum[tid] = sum; s_max[tid] = mx;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {   // inclusive scan of the per-thread step sums, running maximum
    const int a = tid >= d ? s_sum[tid - d] : 0, b = tid >= d ? s_max[tid - d] : 0;
    __syncthreads();
    s_sum[tid] += a; s_max[tid] = max(s_max[tid], b);
    __syncthreads();
  }
  int run = s_sum[tid] - sum;
  for (int n = n0; n < n1; ++n) {
    cs[n] = run;
    run += (p.cu[n + 1] - p.cu[n] + GC - 1) / GC;
  }
  if (tid == 1023) cs[N] = s_sum[1023];
  const int chains = N * Hs, total = Hs * s_sum[1023], longest = s_max[1023];
  const int R = (total + ncta - 1) / ncta;
  const bool cut = chains > ncta && total > 0 && R >= longest;
  if (tid == 0) { p.plan[0] = cut ? 1 : 0; p.plan[1] = ncta; p.plan[2] = 0; p.plan[3] = R; }
  for (int j = tid; j < ncta; j += 1024) p.flags[j] = 0;
  __syncthreads();
  if (!cut) {
    // nothing to balance by cutting: whole chains, dealt round-robin (CTA j: chains j, j + ncta, ...)
    if (tid < ncta) {
      int off = 0, steps = 0;
      for (int j = 0; j < tid; ++j) off += j < chains ? (chains - j + ncta - 1) / ncta : 0;
      int k = 0;
      for (int c = tid; c < chains; c += ncta, ++k) {
        const int n = c / Hs, slot = p.slot[n], mode = p.init_mode[n];
        GlaPiece pc;
        pc.tok_begin = p.cu[n]; pc.len = p.cu[n + 1] - p.cu[n]; pc.hs = c - n * Hs;
        pc.in_kind = 0; pc.in_idx = 0;
        if (mode == 1 && slot >= 0) { pc.in_kind = 1; pc.in_idx = slot; }
        else if (mode == 2 && p.meta_state) pc.in_kind = 2;
        pc.out_kind = 0; pc.out_idx = 0;
        if (p.write_state && slot >= 0) { pc.out_kind = 1; pc.out_idx = slot; }
        pc.pad = 0;
        pieces[off + k] = pc;
        steps += cs[n + 1] - cs[n];
      }
      cta_off[tid] = off; cta_steps[tid] = steps;
      if (tid == ncta - 1) cta_off[ncta] = off + k;
    }
    return;
  }
  // thread j plans range [j R, (j+1) R): pass 0 counts its pieces, pass 1 writes them back to front
  // (an empty chain is a point of the line: it goes to the range that holds the point and only
  // moves its start state to its slot)
  const int lo = tid * R, hi = tid == ncta - 1 ? 0x7fffffff : lo + R;
  int c0 = 0;
  if (tid < ncta) {
    int a = 0, b = chains;             // first chain whose end is >= lo
    while (a < b) {
      const int m = (a + b) >> 1, n = m / Hs, sn = cs[n + 1] - cs[n];
      if (Hs * cs[n] + (m - n * Hs + 1) * sn >= lo) b = m; else a = m + 1;
    }
    c0 = a;
  }
  int off = 0;
  for (int pass = 0; pass < 2; ++pass) {
    int cnt = 0;
    if (tid < ncta) {
      for (int c = c0; c < chains; ++c) {
        const int n = c / Hs, hsx = c - n * Hs, sn = cs[n + 1] - cs[n];
        const int P = Hs * cs[n] + hsx * sn, E = P + sn;
        if (P >= hi) break;
        if (sn > 0 ? E <= lo : P < lo) continue;
        if (pass) {
          const int a = max(P, lo), b = min(E, hi);
          const bool starts = a == P, ends = b == E;
          const int slot = p.slot[n], mode = p.init_mode[n], seq_len = p.cu[n + 1] - p.cu[n];
          GlaPiece pc;
          pc.tok_begin = p.cu[n] + (a - P) * GC;
          pc.len = ends ? seq_len - (a - P) * GC : (b - a) * GC;
          pc.hs = hsx;
          pc.in_kind = 0; pc.in_idx = 0;
          if (!starts) { pc.in_kind = 3; pc.in_idx = tid - 1; }
          else if (mode == 1 && slot >= 0) { pc.in_kind = 1; pc.in_idx = slot; }
          else if (mode == 2 && p.meta_state) pc.in_kind = 2;
          pc.out_kind = 0; pc.out_idx = 0;
          if (!ends) { pc.out_kind = 2; pc.out_idx = tid; }
          else if (p.write_state && slot >= 0) { pc.out_kind = 1; pc.out_idx = slot; }
          pc.pad = 0;
          pieces[off + s_cnt[tid] - 1 - cnt] = pc;
        }
        ++cnt;
      }
    }
    if (pass == 0) {
      s_cnt[tid] = cnt;
      __syncthreads();
      for (int j = 0; j < tid && j < ncta; ++j) off += s_cnt[j];
      if (tid < ncta) {
        cta_off[tid] = off;
        cta_steps[tid] = max(0, min(hi, total) - min(lo, total));
        if (tid == ncta - 1) cta_off[ncta] = off + cnt;
      }
    }
  }
}

}  // namespace hymeta

using namespace hymeta;

#ifdef HYMETA_TRACE
extern "C" int hymeta_debug_set_gla_trace(void* buf) {
  return (int)cudaMemcpyToSymbol(g_gla_trace, &buf, sizeof(buf));
}
#endif

extern "C" size_t hymeta_gla_prefill_workspace_bytes(int N, int Hs, int max_seqlen, int seg_len) {
  if (seg_len <= 0 || N <= 0 || Hs <= 0 || max_seqlen <= seg_len) return 0;
  const size_t n_segs = ((size_t)max_seqlen + seg_len - 1) / seg_len;
  return (size_t)N * Hs * n_segs * GP_WS_ELEMS * sizeof(float);
}

// Balanced single-pass form (seg_len == HYMETA_GLA_SEG_BALANCED): workspace = one fp32 hand-off
// buffer per SM, then the plan (header, per-CTA ranges, step totals, flags, step prefix, pieces).
static size_t balanced_ints(int N, int ncta) {
  return ((size_t)GP_PLAN_HDR + (ncta + 1) + ncta + ncta + (N + 1) + 3) & ~(size_t)3;
}
extern "C" size_t hymeta_gla_prefill_balanced_workspace_bytes(int N, int Hs) {
  if (N <= 0 || Hs <= 0) return 0;
  const int ncta = sm_count();
  return (size_t)ncta * GP_WS_ELEMS * sizeof(float) + balanced_ints(N, ncta) * sizeof(int32_t) +
         ((size_t)N * Hs + ncta) * sizeof(GlaPiece);
}

extern "C" int hymeta_gla_prefill_varlen_fused(
    const void* q, const void* k, const void* v, const void* g, void* o, void* state,
    const void* meta_state, const int32_t* cu_seqlens, const int32_t* slot_idx,
    const int32_t* init_mode, int N, int T, int H, int Hk, int Hv, int Hs, int D, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t g_st, int64_t g_sh, int64_t o_st,
    int64_t o_sh, int64_t st_ss, int64_t st_sh, int64_t meta_sh, int g_is_fp32, float scale,
    int write_state, int max_seqlen, int seg_len, void* workspace, size_t workspace_bytes,
    const void* gnorm_weight, float gnorm_eps, int relu_qk, const void* gate_table, float* sumsq_out,
    void* stream) {
  if (N == 0) return HYMETA_OK;
  if (gate_table && g_is_fp32) return HYMETA_ERR_INVALID_ARG;   // the table maps bf16 gate bits
  const bool balanced = seg_len == HYMETA_GLA_SEG_BALANCED;
  if (balanced) {
    if (N < 0 || Hs <= 0) return HYMETA_ERR_INVALID_ARG;
    if (!workspace || ((uintptr_t)workspace % 16)) return HYMETA_ERR_INVALID_ARG;
    if (workspace_bytes < hymeta_gla_prefill_balanced_workspace_bytes(N, Hs)) return HYMETA_ERR_WORKSPACE;
    seg_len = 0;
  }
  if (seg_len < 0 || seg_len % GC) return HYMETA_ERR_INVALID_ARG;
  int n_segs = 1;
  if (seg_len > 0 && max_seqlen > seg_len) {
    n_segs = (max_seqlen + seg_len - 1) / seg_len;
    if (n_segs > 65535) return HYMETA_ERR_UNSUPPORTED;
    if (!workspace || workspace_bytes < hymeta_gla_prefill_workspace_bytes(N, Hs, max_seqlen, seg_len) ||
        ((uintptr_t)workspace % 16))
      return HYMETA_ERR_INVALID_ARG;
  } else seg_len = 0;
  if (N < 0 || T < 0 || H <= 0 || Hk <= 0 || Hv <= 0 || Hs <= 0 || !cu_seqlens || !slot_idx || !init_mode || !state)
    return HYMETA_ERR_INVALID_ARG;
  if (T > 0 && (!q || !k || !v || !g || !o)) return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  if (H % Hs || Hs % Hk || Hs % Hv) return HYMETA_ERR_INVALID_ARG;
  if (H / Hs > 8) return HYMETA_ERR_UNSUPPORTED;   // q-heads per state are stacked into one M=128 tile
  if (N > 65535) return HYMETA_ERR_UNSUPPORTED;
  const int64_t vec[] = {q_st, q_sh, v_st, v_sh, o_st, o_sh, st_ss, st_sh, meta_state ? meta_sh : 0};
  for (int64_t s : vec)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;
  GlaPrefillParams p;
  p.q = (const __nv_bfloat16*)q; p.k = (const __nv_bfloat16*)k; p.v = (const __nv_bfloat16*)v; p.g = g;
  p.o = (__nv_bfloat16*)o; p.state = (__nv_bfloat16*)state; p.meta_state = (const __nv_bfloat16*)meta_state;
  p.cu = cu_seqlens; p.slot = slot_idx; p.init_mode = init_mode;
  p.N = N; p.H = H; p.Hk = Hk; p.Hv = Hv; p.Hs = Hs; p.write_state = write_state; p.scale = scale;
  p.q_st = q_st; p.q_sh = q_sh; p.k_st = k_st; p.k_sh = k_sh; p.v_st = v_st; p.v_sh = v_sh;
  p.g_st = g_st; p.g_sh = g_sh; p.o_st = o_st; p.o_sh = o_sh; p.st_ss = st_ss; p.st_sh = st_sh;
  p.meta_sh = meta_sh;
  if (gnorm_weight && ((uintptr_t)gnorm_weight & 15)) return HYMETA_ERR_INVALID_ARG;
  p.gnorm_w = (const __nv_bfloat16*)gnorm_weight; p.gnorm_eps = gnorm_eps;
  p.relu_lim = relu_qk ? 0u : 0xff80ff80u;
  p.gate_table = (const uint16_t*)gate_table;
  p.sumsq = sumsq_out;
  p.seg_len = seg_len; p.n_segs = n_segs; p.pass = 0; p.ws = (float*)workspace;
  p.plan = nullptr; p.cta_off = p.cta_steps = nullptr; p.pieces = nullptr; p.flags = nullptr;
  int32_t* plan_cs = nullptr;
  const int ncta = sm_count();
  if (balanced) {
    int32_t* ints = reinterpret_cast<int32_t*>((float*)workspace + (size_t)ncta * GP_WS_ELEMS);
    p.plan = ints;
    p.cta_off = ints + GP_PLAN_HDR;
    p.cta_steps = ints + GP_PLAN_HDR + ncta + 1;
    p.flags = ints + GP_PLAN_HDR + 2 * ncta + 1;
    plan_cs = ints + GP_PLAN_HDR + 3 * ncta + 1;
    p.pieces = reinterpret_cast<const GlaPiece*>(ints + balanced_ints(N, ncta));
  }
  // TMA maps: raw g / k rows ([16 tokens][128 channels] boxes) and v (SWIZZLE_128B 64-wide halves)
  const int64_t tvec[] = {k_st, k_sh, g_st * (g_is_fp32 ? 2 : 1), g_sh * (g_is_fp32 ? 2 : 1)};
  for (int64_t s : tvec)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;   // TMA needs 16-byte multiples
  CUtensorMap tmG, tmK, tmV;
  memset(&tmG, 0, sizeof(tmG)); tmK = tmV = tmG;
  if (T > 0) {   // with no tokens every chain is empty and the maps are never dereferenced
    const uint64_t dk[3] = {(uint64_t)kD, (uint64_t)Hk, (uint64_t)T};
    const uint64_t dv[3] = {(uint64_t)kD, (uint64_t)Hv, (uint64_t)T};
    const int64_t sg[3] = {1, g_sh, g_st}, sk[3] = {1, k_sh, k_st}, sv[3] = {1, v_sh, v_st};
    const uint32_t braw[3] = {(uint32_t)kD, 1, (uint32_t)GC}, bv[3] = {64, 1, (uint32_t)GC};
    int rc;
    if ((rc = make_tmap(&tmG, g, g_is_fp32 ? 4 : 2, 3, dk, sg, braw, false))) return rc;
    if ((rc = make_tmap(&tmK, k, 2, 3, dk, sk, braw, false))) return rc;
    if ((rc = make_tmap(&tmV, v, 2, 3, dv, sv, bv, true))) return rc;
  }
  static DynSmemAttr attr;
  if (attr.need()) {
    DynSmemAttr::set(gla_prefill_kernel<true, false>, GP_SMEM);
    DynSmemAttr::set(gla_prefill_kernel<false, false>, GP_SMEM);
    DynSmemAttr::set(gla_prefill_kernel<true, true>, GP_SMEM);
    DynSmemAttr::set(gla_prefill_kernel<false, true>, GP_SMEM);
    attr.mark();
  }
  auto launch = [&](int pass, int segs) {
    p.pass = pass;
    dim3 grid(Hs, N, segs);
    if (balanced) grid = dim3(ncta, 1, 1);
    cudaStream_t st = (cudaStream_t)stream;
    const bool gn = p.gnorm_w != nullptr && pass != 1;   // the summary pass emits no outputs
    if (g_is_fp32) {
      if (gn) gla_prefill_kernel<true, true><<<grid, GP_THREADS, GP_SMEM, st>>>(p, tmG, tmK, tmV);
      else gla_prefill_kernel<true, false><<<grid, GP_THREADS, GP_SMEM, st>>>(p, tmG, tmK, tmV);
    } else {
      if (gn) gla_prefill_kernel<false, true><<<grid, GP_THREADS, GP_SMEM, st>>>(p, tmG, tmK, tmV);
      else gla_prefill_kernel<false, false><<<grid, GP_THREADS, GP_SMEM, st>>>(p, tmG, tmK, tmV);
    }
    count_launch();
  };
  if (balanced) {
    gla_prefill_plan_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(
        p, ncta, plan_cs, const_cast<int32_t*>(p.cta_off), const_cast<int32_t*>(p.cta_steps),
        const_cast<GlaPiece*>(p.pieces));
    count_launch();
    launch(0, 1);                                  // one persistent CTA per SM, whatever the plan says
  } else if (n_segs == 1) launch(0, 1);
  else {
    // (1) per-segment summaries from a zero start, (2) scan into segment start states, (3) outputs
    launch(1, n_segs - 1);
    gla_prefill_scan_kernel<<<dim3(Hs, N, kD * kD / 1024), 256, 0, (cudaStream_t)stream>>>(p);
    count_launch();
    launch(3, n_segs);
  }
  return check_launch();
}

extern "C" int hymeta_gla_prefill_varlen_seg_gnorm(
    const void* q, const void* k, const void* v, const void* g, void* o, void* state,
    const void* meta_state, const int32_t* cu_seqlens, const int32_t* slot_idx,
    const int32_t* init_mode, int N, int T, int H, int Hk, int Hv, int Hs, int D, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t g_st, int64_t g_sh, int64_t o_st,
    int64_t o_sh, int64_t st_ss, int64_t st_sh, int64_t meta_sh, int g_is_fp32, float scale,
    int write_state, int max_seqlen, int seg_len, void* workspace, size_t workspace_bytes,
    const void* gnorm_weight, float gnorm_eps, void* stream) {
  return hymeta_gla_prefill_varlen_fused(q, k, v, g, o, state, meta_state, cu_seqlens, slot_idx, init_mode, N, T,
                                         H, Hk, Hv, Hs, D, q_st, q_sh, k_st, k_sh, v_st, v_sh, g_st, g_sh,
                                         o_st, o_sh, st_ss, st_sh, meta_sh, g_is_fp32, scale, write_state,
                                         max_seqlen, seg_len, workspace, workspace_bytes, gnorm_weight,
                                         gnorm_eps, 0, nullptr, nullptr, stream);
}

extern "C" int hymeta_gla_prefill_varlen_seg(
    const void* q, const void* k, const void* v, const void* g, void* o, void* state,
    const void* meta_state, const int32_t* cu_seqlens, const int32_t* slot_idx,
    const int32_t* init_mode, int N, int T, int H, int Hk, int Hv, int Hs, int D, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t g_st, int64_t g_sh, int64_t o_st,
    int64_t o_sh, int64_t st_ss, int64_t st_sh, int64_t meta_sh, int g_is_fp32, float scale,
    int write_state, int max_seqlen, int seg_len, void* workspace, size_t workspace_bytes, void* stream) {
  return hymeta_gla_prefill_varlen_seg_gnorm(q, k, v, g, o, state, meta_state, cu_seqlens, slot_idx, init_mode,
                                             N, T, H, Hk, Hv, Hs, D, q_st, q_sh, k_st, k_sh, v_st, v_sh, g_st,
                                             g_sh, o_st, o_sh, st_ss, st_sh, meta_sh, g_is_fp32, scale,
                                             write_state, max_seqlen, seg_len, workspace, workspace_bytes,
                                             nullptr, 0.f, stream);
}

extern "C" int hymeta_gla_prefill_varlen(
    const void* q, const void* k, const void* v, const void* g, void* o, void* state,
    const void* meta_state, const int32_t* cu_seqlens, const int32_t* slot_idx,
    const int32_t* init_mode, int N, int T, int H, int Hk, int Hv, int Hs, int D, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t g_st, int64_t g_sh, int64_t o_st,
    int64_t o_sh, int64_t st_ss, int64_t st_sh, int64_t meta_sh, int g_is_fp32, float scale,
    int write_state, int max_seqlen, void* stream) {
  return hymeta_gla_prefill_varlen_seg(q, k, v, g, o, state, meta_state, cu_seqlens, slot_idx, init_mode, N, T,
                                       H, Hk, Hv, Hs, D, q_st, q_sh, k_st, k_sh, v_st, v_sh, g_st, g_sh, o_st,
                                       o_sh, st_ss, st_sh, meta_sh, g_is_fp32, scale, write_state, max_seqlen,
                                       0, nullptr, 0, stream);
}
