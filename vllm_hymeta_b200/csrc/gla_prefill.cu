// GLA chunked prefill on tcgen05 / TMEM (a1/a3), varlen over cu_seqlens with slot
// states -- replaces the reference's per-request Python loop around fused_chunk_gla, its two host
// syncs per request and the 7x repeat_kv materialisation (modeling_hymeta.py:441-549).
//
// One CTA (one per SM, all 512 TMEM columns) walks one (sequence, state-head[, segment]) chain in
// steps of C = 16 tokens.  The G = H/Hs q-heads sharing the state are stacked along the MMA M
// dimension: row = head*16 + token (G <= 8 -> M = 128).  With b_t = cumsum(g) inside the step and
// the per-channel reference r = b_7 (mid-step, so every exponent is bounded by 8 steps of decay):
//     q~_t = q_t * scale * exp(b_t - r)      k~_s = k_s * exp(r - b_s)
//     A    = q~ k~^T  (causal, s <= t)               MMA  M128 N16  K128   (K-major x K-major)
//     O    = q~ X  +  P v      (P = bf16 causal A)    MMA  M128 N128 K128 (SS) + K16 (TS, P in TMEM)
//     Y    = X + k~^T v                               MMA  M128 N128 K16    (MN-major x MN-major)
//     X'   = exp(b_15 - r + r') (.) Y                 fp32 master in TMEM, rescaled in place
// where X = exp(r) (.) S is the state expressed in the step's reference frame.  The fp32 master
// never leaves TMEM; a bf16 copy (the MN-major B operand of the next step's q~ X) is rewritten to
// shared memory by the same pass that rescales it.
//
// Warp roles (544 threads), all hand-offs through mbarriers:
//   warps 0-3   gates: thread = channel; loads g/k/v of a step, builds k~, the q-side factor table
//               scale*exp(b_t - r) and the per-channel rescale factors into a 4-stage ring
//   warps 4-7   state: thread = d_k row = TMEM lane; X' = D (.) Y, bf16 copy, final state store
//   warps 8-11  q-tile builder / P converter / O epilogue for even steps (thread = row for TMEM)
//   warps 12-15 same for odd steps
//   warp 16     MMA issuer (one lane)
#include "umma.cuh"

namespace hymeta {

constexpr int GC = 16;                          // tokens per step
constexpr int GP_KS = 4;                        // gate ring stages
constexpr int GP_SQ = 0;                        // q~  2 x [2][128][64] bf16            64 KB
constexpr int GP_SS = 65536;                    // X   2 x [2][128 dk][64 dv] bf16      64 KB
constexpr int GP_ST = 131072;                   // gate ring
constexpr int GP_ST_K = 0;                      //   k~ [2][16][64] bf16                 4 KB
constexpr int GP_ST_V = 4096;                   //   v  [2][16][64] bf16                 4 KB
constexpr int GP_ST_F = 8192;                   //   scale*exp(b_t - r) [16][128] f32    8 KB
constexpr int GP_ST_D = 16384;                  //   exp(b_15 - r) [128], exp(r) [128]   1 KB
constexpr int GP_STAGE = 17408;
constexpr int GP_BAR = GP_ST + GP_KS * GP_STAGE;
constexpr int GP_SMEM = GP_BAR + 256;
constexpr int GP_THREADS = 512;

struct GlaPrefillParams {
  const __nv_bfloat16 *q, *k, *v;
  const void* g;
  __nv_bfloat16 *o, *state;
  const __nv_bfloat16* meta_state;
  const int32_t *cu, *slot, *init_mode;
  int N, H, Hk, Hv, Hs, write_state;
  float scale;
  int64_t q_st, q_sh, k_st, k_sh, v_st, v_sh, g_st, g_sh, o_st, o_sh, st_ss, st_sh, meta_sh;
};

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
#ifndef GP_EXP
#define GP_EXP 0   // timing experiments of tools/trace_gla_prefill.py (bit mask); 0 in the product build
#endif

template <bool GF32>
__global__ void __launch_bounds__(GP_THREADS, 1) gla_prefill_kernel(const GlaPrefillParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int hs = blockIdx.x, n = blockIdx.y;
  const int G = p.H / p.Hs, h0 = hs * G, hk = hs / (p.Hs / p.Hk), hv = hs / (p.Hs / p.Hv);
  const int tok_begin = p.cu[n], len = p.cu[n + 1] - tok_begin;
  const int slot = p.slot[n], mode = p.init_mode[n];
  const int n_steps = (len + GC - 1) / GC;

  const __nv_bfloat16* init_src = nullptr;
  if (mode == 1 && slot >= 0) init_src = p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh;
  else if (mode == 2 && p.meta_state) init_src = p.meta_state + (int64_t)hs * p.meta_sh;
  __nv_bfloat16* final_dst =
      (p.write_state && slot >= 0) ? p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh : nullptr;

  if (n_steps == 0) {   // empty sequence: the slot receives the initial state unchanged
    if (final_dst && final_dst != init_src && warp >= 4 && warp < 8) {
      const int d = tid - 128;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        uint4 t = make_uint4(0, 0, 0, 0);
        if (init_src) t = ldg_v4(init_src + d * kD + 8 * c);
        stg_v4(final_dst + d * kD + 8 * c, t);
      }
    }
    return;
  }

  const uint32_t sbase = smem_u32(smem);
  const uint32_t sQ = sbase + GP_SQ, sS = sbase + GP_SS, sT = sbase + GP_ST;
  const uint32_t bars = sbase + GP_BAR;
  const uint32_t bar_kf = bars, bar_ke = bar_kf + 8 * GP_KS, bar_qf = bar_ke + 8 * GP_KS,
                 bar_a = bar_qf + 16, bar_p = bar_a + 16, bar_o = bar_p + 16, bar_x = bar_o + 16,
                 bar_s = bar_x + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + GP_BAR + 200);

  if (tid == 0) {
    if (sbase & 1023) asm volatile("trap;");
    for (int s = 0; s < GP_KS; ++s) { mbar_init(bar_kf + 8 * s, 96); mbar_init(bar_ke + 8 * s, 257); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_qf + 8 * i, 128); mbar_init(bar_a + 8 * i, 1);
      mbar_init(bar_p + 8 * i, 128);  mbar_init(bar_o + 8 * i, 1);
    }
    mbar_init(bar_x, 128); mbar_init(bar_s, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(smem_u32(tmem_slot));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // TMEM columns: X/Y [0,128)  O0 [128,256)  O1 [256,384)  A0 [384,400)  A1 [400,416); P_i aliases A_i
  const uint32_t tX = tmem, tO = tmem + 128, tA = tmem + 384;

  if (warp < 2) {
    // ================================================================== gates (thread = 2 channels)
    // Straight-line code is kept small on purpose: five roles share the SM's instruction cache.
    const int c = 2 * tid;
    constexpr int GB = GF32 ? 4 : 2;
    const char* gp = reinterpret_cast<const char*>(p.g) + ((int64_t)tok_begin * p.g_st + (int64_t)hk * p.g_sh + c) * GB;
    const char* kp = reinterpret_cast<const char*>(p.k + (int64_t)tok_begin * p.k_st + (int64_t)hk * p.k_sh + c);
    const int64_t gstep = p.g_st * GB, kstep = p.k_st * 2;
    const uint32_t kbase = (uint32_t)((c >> 6) * 2048 + (c & 7) * 2), kcx = (uint32_t)((c >> 3) & 7);
    const float lscale = __log2f(p.scale);
    float2 g[GC];
    uint32_t kk[GC];
    for (int i = -1; i < n_steps; ++i) {
      if (i >= 0) {
        const int st = i % GP_KS, clen = len - i * GC;
        uint8_t* stage = smem + GP_ST + st * GP_STAGE;
        // cumulative gates in the log2 domain; padded tokens (t >= clen) do not decay
        float2 b[GC];
        float bx = 0.f, by = 0.f;
#pragma unroll
        for (int t = 0; t < GC; ++t) {
          if (t < clen) { bx = fmaf(g[t].x, 1.4426950408889634f, bx); by = fmaf(g[t].y, 1.4426950408889634f, by); }
          b[t] = make_float2(bx, by);
        }
        const float2 r = b[GC / 2 - 1];
        const float rfx = r.x - lscale, rfy = r.y - lscale;
        if (i >= GP_KS) mbar_wait(bar_ke + 8 * st, ((i / GP_KS) - 1) & 1);
        uint8_t* sF = stage + GP_ST_F + c * 4;
        uint8_t* sK = stage + GP_ST_K + kbase;
#pragma unroll
        for (int t = 0; t < GC; ++t) {
          *reinterpret_cast<float2*>(sF + t * (kD * 4)) = make_float2(exp2f(b[t].x - rfx), exp2f(b[t].y - rfy));
          *reinterpret_cast<uint32_t*>(sK + t * 128 + ((kcx ^ (t & 7)) << 4)) =
              pack_bf16(bf16_lo(kk[t]) * exp2f(r.x - b[t].x), bf16_hi(kk[t]) * exp2f(r.y - b[t].y));
        }
        float* sD = reinterpret_cast<float*>(stage + GP_ST_D);
        *reinterpret_cast<float2*>(sD + c) = make_float2(exp2f(b[GC - 1].x - r.x), exp2f(b[GC - 1].y - r.y));
        *reinterpret_cast<float2*>(sD + kD + c) = make_float2(exp2f(r.x), exp2f(r.y));
        fence_proxy_async_smem();
        mbar_arrive(bar_kf + 8 * st);
        if (tid == 0) GP_TRACE(i, 0);
      }
      if (i + 1 < n_steps) {   // next step's raw gates / keys: rows past the sequence end re-read its last token
        const int base = (i + 1) * GC;
#pragma unroll
        for (int t = 0; t < GC; ++t) {
          const int64_t tt = min(base + t, len - 1);
          if (GF32) g[t] = *reinterpret_cast<const float2*>(gp + tt * gstep);
          else {
            const uint32_t u = *reinterpret_cast<const uint32_t*>(gp + tt * gstep);
            g[t] = make_float2(bf16_lo(u), bf16_hi(u));
          }
          kk[t] = *reinterpret_cast<const uint32_t*>(kp + tt * kstep);
        }
      }
    }
  } else if (warp == 3) {
    // ================================================================== v loader (one warp)
    const int vc = lane & 15, vt = lane >> 4;       // 16-byte chunk vc of tokens vt + 2j
    const __nv_bfloat16* vp = p.v + (int64_t)(tok_begin + vt) * p.v_st + (int64_t)hv * p.v_sh + vc * 8;
    const uint32_t voff = (uint32_t)(GP_ST_V + (vc >> 3) * 2048 + vt * 128);
    uint4 x[8];
    for (int i = -1; i < n_steps; ++i) {
      if (i >= 0) {
        const int st = i % GP_KS;
        uint8_t* sv = smem + GP_ST + st * GP_STAGE + voff;
        if (i >= GP_KS) mbar_wait(bar_ke + 8 * st, ((i / GP_KS) - 1) & 1);
#pragma unroll
        for (int j = 0; j < 8; ++j)   // row vt + 2j: 16-byte slot (vc ^ row) & 7
          *reinterpret_cast<uint4*>(sv + j * 256 + (((vc ^ (vt + 2 * j)) & 7) << 4)) = x[j];
        fence_proxy_async_smem();
        mbar_arrive(bar_kf + 8 * st);
        if (lane == 0) GP_TRACE(i, 1);
      }
      if (i + 1 < n_steps) {
        const int rem = len - (i + 1) * GC - vt;    // tokens vt + 2j < clen
        const __nv_bfloat16* src = vp + (int64_t)(i + 1) * GC * p.v_st;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          x[j] = make_uint4(0, 0, 0, 0);
          if (2 * j < rem) x[j] = ldg_nc_v4(src + (int64_t)(2 * j) * p.v_st);
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // ================================================================== state (thread = d_k row)
    // pass i = -1 loads S_init and scales it into the first step's frame; pass i >= 0 rescales Y_i
    // into the next frame; the last pass stores the final state.  One copy of the 32-column body.
    const int d = tid - 128;
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const float* sD0 = reinterpret_cast<const float*>(smem + GP_ST + GP_ST_D) + d;
    const __nv_bfloat16* src = init_src ? init_src + d * kD : nullptr;
    __nv_bfloat16* dst = final_dst ? final_dst + d * kD : nullptr;
    const uint32_t srow = (uint32_t)(GP_SS + d * 128), sx = (uint32_t)(d & 7);
    for (int i = -1; i < n_steps; ++i) {
      const bool first = i < 0, last = i + 1 == n_steps;
      float D = first ? 1.f : sD0[(i % GP_KS) * (GP_STAGE / 4)];
      if (!last) {
        const int sn = (i + 1) % GP_KS;
        mbar_wait(bar_kf + 8 * sn, ((i + 1) / GP_KS) & 1);
        D *= sD0[sn * (GP_STAGE / 4) + kD];
      }
      if (!first) {
        mbar_wait(bar_s, i & 1);   // Y_i complete; every earlier MMA (incl. q~ X of step i-1) too
        tc_fence_after();
        if (d == 0) GP_TRACE(i, 2);
      }
      if (!last || dst) {
        uint8_t* sdst = smem + srow + ((i + 1) & 1) * 32768;
#pragma unroll 1
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t x[32];
          if (first) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 t = make_uint4(0, 0, 0, 0);
              if (src) t = ldg_v4(src + 32 * cc + 8 * j);
              x[8 * j + 0] = t.x << 16; x[8 * j + 1] = t.x & 0xffff0000u; x[8 * j + 2] = t.y << 16;
              x[8 * j + 3] = t.y & 0xffff0000u; x[8 * j + 4] = t.z << 16; x[8 * j + 5] = t.z & 0xffff0000u;
              x[8 * j + 6] = t.w << 16; x[8 * j + 7] = t.w & 0xffff0000u;
            }
          } else {
            tmem_ld32_nowait(tX + lane_off + 32 * cc, x);
            tmem_wait_ld();
            if (d == 0 && cc == 0) GP_TRACE(i, 14);
          }
#pragma unroll
          for (int u = 0; u < 32; ++u) x[u] = __float_as_uint(__uint_as_float(x[u]) * D);
          if (!last) tmem_st32(tX + lane_off + 32 * cc, x);
          if (d == 0 && cc == 0 && i >= 0) GP_TRACE(i, 15);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 w4;
            w4.x = pack_bf16(__uint_as_float(x[8 * j + 0]), __uint_as_float(x[8 * j + 1]));
            w4.y = pack_bf16(__uint_as_float(x[8 * j + 2]), __uint_as_float(x[8 * j + 3]));
            w4.z = pack_bf16(__uint_as_float(x[8 * j + 4]), __uint_as_float(x[8 * j + 5]));
            w4.w = pack_bf16(__uint_as_float(x[8 * j + 6]), __uint_as_float(x[8 * j + 7]));
            const uint32_t ch = (uint32_t)(4 * cc + j);
            if (!last) { if (!(GP_EXP & 2)) *reinterpret_cast<uint4*>(sdst + (ch >> 3) * 16384 + (((ch ^ sx) & 7) << 4)) = w4; }
            else stg_v4(dst + 32 * cc + 8 * j, w4);
          }
        }
      }
      if (!last) {
        tmem_wait_st();
        fence_proxy_async_smem();
        tc_fence_before();
        mbar_arrive(bar_x);
        if (d == 0 && i >= 0) GP_TRACE(i, 3);
      }
      if (!first) mbar_arrive(bar_ke + 8 * (i % GP_KS));
    }
  } else if (warp >= 8) {
    // ================================================================== q tiles (even / odd steps)
    const int w = (warp - 8) >> 2;
    const int tq = tid - 256 - w * 128;             // 0..127
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const int qc = tq & 15, qt0 = tq >> 4;          // builder view: 16-byte chunk qc, tokens qt0, qt0+8
    const int r_hh = tq >> 4, r_t = tq & 15;        // row-owner view: head slot, token
    const uint32_t tAw = tA + 16 * w + lane_off, tOw = tO + 128 * w + lane_off;
    // q~ row (hh*16 + t) of 16-byte chunk qc: byte offsets of tokens qt0 / qt0+8 for head slot 0
    uint8_t* sQw = smem + GP_SQ + w * 32768 + (qc >> 3) * 16384;
    const uint32_t qo0 = (uint32_t)(qt0 * 128 + (((qc ^ qt0) & 7) << 4)), qo1 = qo0 + 1024;   // (qt0+8)&7 == qt0&7
    const __nv_bfloat16* qp = p.q + (int64_t)(tok_begin + qt0) * p.q_st + (int64_t)h0 * p.q_sh + qc * 8;
    __nv_bfloat16* op = p.o + (int64_t)(tok_begin + r_t) * p.o_st + (int64_t)(h0 + r_hh) * p.o_sh;
    uint4 qv[16];
    int cnt = 0;
    for (int i = w - 2; i < n_steps; i += 2) {
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
          o4.x = pack_bf16(bf16_lo(v4.x) * e[0], bf16_hi(v4.x) * e[1]);
          o4.y = pack_bf16(bf16_lo(v4.y) * e[2], bf16_hi(v4.y) * e[3]);
          o4.z = pack_bf16(bf16_lo(v4.z) * e[4], bf16_hi(v4.z) * e[5]);
          o4.w = pack_bf16(bf16_lo(v4.w) * e[6], bf16_hi(v4.w) * e[7]);
          if (!(GP_EXP & 4)) *reinterpret_cast<uint4*>(sQw + (it >> 1) * 2048 + ((it & 1) ? qo1 : qo0)) = o4;
        }
        fence_proxy_async_smem();
        tc_fence_before();                          // orders the previous epilogue's TMEM reads too
        mbar_arrive(bar_qf + 8 * w);
        if (tq == 0) GP_TRACE(i, 5);
      }
      if (i + 2 < n_steps) {                        // next tile: in flight under the P conversion and epilogue
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
          const bool live = r_hh < G && r_t < clen;
          __nv_bfloat16* dst = op + (int64_t)i * GC * p.o_st;
#pragma unroll 1
          for (int cc = 0; cc < ((GP_EXP & 1) ? 0 : 4); ++cc) {
            float t[32];
            tmem_ld32(tOw + 32 * cc, t);
            if (live) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint4 o4;
                o4.x = pack_bf16(t[8 * j + 0], t[8 * j + 1]); o4.y = pack_bf16(t[8 * j + 2], t[8 * j + 3]);
                o4.z = pack_bf16(t[8 * j + 4], t[8 * j + 5]); o4.w = pack_bf16(t[8 * j + 6], t[8 * j + 7]);
                stg_v4(dst + 32 * cc + 8 * j, o4);
              }
            }
          }
        }
        ++cnt;
      }
    }
  } else if (warp == 2) {
    // ================================================================== MMA issuer
    if (lane == 0) {
      const uint32_t idesc_a = umma_idesc_bf16(128, 16, 0, 0);
      const uint32_t idesc_o = umma_idesc_bf16(128, 128, 0, 1);
      const uint32_t idesc_s = umma_idesc_bf16(128, 128, 1, 1);
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
        const uint32_t kb = (sT + st * GP_STAGE + GP_ST_K) >> 4, vb = (sT + st * GP_STAGE + GP_ST_V) >> 4;
        GP_TRACE(i, 9);
        mbar_wait(bar_x, i & 1);
        tc_fence_after();
        GP_TRACE(i, 10);
        umma_bf16(tX, dmn + kb, dmn + vb, idesc_s, 1);   // Y = X + k~^T v
        umma_commit(bar_s);
        {                                                // O = q~ X
          const uint32_t qb = (sQ + w * 32768) >> 4, xb = (sS + w * 32768) >> 4;
#pragma unroll 1
          for (int ks = 0; ks < 8; ++ks)
            umma_bf16(tO + 128 * w, dk + qb + (ks >> 2) * 1024 + (ks & 3) * 2, dx + xb + ks * 128, idesc_o, ks > 0);
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
        if (!a_done) {
          mbar_wait(bar_kf + 8 * ((i + 1) % GP_KS), ((i + 1) / GP_KS) & 1);
          mbar_wait(bar_qf + 8 * (w ^ 1), ((i + 1) >> 1) & 1);
          tc_fence_after();
          issue_a(i + 1);
        }
        umma_commit(bar_ke + 8 * st);
        GP_TRACE(i, 13);
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace hymeta

using namespace hymeta;

#ifdef HYMETA_TRACE
extern "C" int hymeta_debug_set_gla_trace(void* buf) {
  return (int)cudaMemcpyToSymbol(g_gla_trace, &buf, sizeof(buf));
}
#endif

extern "C" int hymeta_gla_prefill_varlen(
    const void* q, const void* k, const void* v, const void* g, void* o, void* state,
    const void* meta_state, const int32_t* cu_seqlens, const int32_t* slot_idx,
    const int32_t* init_mode, int N, int T, int H, int Hk, int Hv, int Hs, int D, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t g_st, int64_t g_sh, int64_t o_st,
    int64_t o_sh, int64_t st_ss, int64_t st_sh, int64_t meta_sh, int g_is_fp32, float scale,
    int write_state, int max_seqlen, void* stream) {
  (void)max_seqlen;
  if (N == 0) return HYMETA_OK;
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
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(gla_prefill_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GP_SMEM);
    cudaFuncSetAttribute(gla_prefill_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GP_SMEM);
    attr_set = true;
  }
  dim3 grid(Hs, N);
  if (g_is_fp32) gla_prefill_kernel<true><<<grid, GP_THREADS, GP_SMEM, (cudaStream_t)stream>>>(p);
  else gla_prefill_kernel<false><<<grid, GP_THREADS, GP_SMEM, (cudaStream_t)stream>>>(p);
  count_launch();
  return check_launch();
}
