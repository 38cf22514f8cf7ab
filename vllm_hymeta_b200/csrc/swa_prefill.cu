// SWA + meta-prefix varlen prefill on tcgen05 / TMEM / TMA (a6/a7).
//
// One launch replaces the reference's two flash-attn launches (windowed causal pass + non-causal
// meta pass), the concat of q, the 6-kernel TorchScript LSE merge and the output concat
// (fa_kernel_metatoken.py:198-285): the meta keys are simply the first key tiles of every query
// tile's single online softmax.
//
// CTA = TWO 128-row query tiles of one q-head (256 consecutive queries) against 128-key tiles,
// warp-specialised, one CTA per SM (all 512 TMEM columns, 224 KB of smem):
//   warp 8      TMA producer: Q0/Q1 once, then K tiles (2-stage ring) and V tiles (3-stage ring),
//               SWIZZLE_128B boxes straight from the packed [T, Hkv, D] tensors
//   warp 9      MMA issuer (one lane):  S_i = Q_i K^T          tcgen05.mma SS, fp32 S in TMEM
//                                       O_i += P_i V           tcgen05.mma TS: A = bf16 P in TMEM
//                                                              (aliases S_i), B = V MN-major as landed
//   warps 0-3   softmax of query tile 0, warps 4-7 of tile 1 (thread = query row = TMEM lane):
//               one tcgen05.ld of the 128 scores, masks only on boundary tiles, exp2 in the log2
//               domain, bf16 P stored back over S with tcgen05.st.  O stays in TMEM for the whole
//               key loop; it is rescaled in place only when a row maximum grows by more than 2^8
//               (lazy rescale), so the common step touches neither O nor shared memory.
// The issue order  PV0(t) S0(t+1) PV1(t) S1(t+1)  lets tile 1's softmax run under tile 0's MMAs and
// vice versa; MMAs of one CTA complete in order, so "S_i(t) has landed" also means O_i holds every
// PV up to t-1 and the rescale needs no extra barrier.
// grid.x = q-head so the G heads of a GQA group run side by side and share K/V tiles through L2.
#include "umma.cuh"

namespace hymeta {

constexpr int PF_BM = 128;    // query rows per tile (two tiles per CTA)
constexpr int PF_BN = 128;    // keys per tile
constexpr int PF_NKS = 2;     // K ring stages
constexpr int PF_NVS = 3;     // V ring stages
constexpr int PF_TILE = 32768;                          // one [128][128] bf16 tile: 2 halves x [128][64]
constexpr int PF_SQ = 0;                                // Q0, Q1
constexpr int PF_SK = 2 * PF_TILE;
constexpr int PF_SV = PF_SK + PF_NKS * PF_TILE;
constexpr int PF_BAR = PF_SV + PF_NVS * PF_TILE;        // mbarriers + tmem ptr
constexpr int PF_SMEM = PF_BAR + 256;
#ifndef PF_OPT_REGS
#define PF_OPT_REGS 1     // 384 threads (3 warpgroups) and setmaxnreg: softmax warps PF_REG_HI registers, the rest PF_REG_LO
#endif
#ifndef PF_OPT_PSPLIT
#define PF_OPT_PSPLIT 2   // P handed to the MMA warp in this many column groups (1, 2 or 4), each with its own mbarrier
#endif
#ifndef PF_REG_LO
#define PF_REG_LO 72
#define PF_REG_HI 216
#endif
constexpr int PF_THREADS = PF_OPT_REGS ? 384 : 320;
constexpr int PF_PS = PF_OPT_PSPLIT;
constexpr float PF_RESCALE_T = 8.f;                     // lazy rescale threshold (log2 domain)

struct SwaPrefillParams {
  const int32_t *cu_q, *cu_k;
  __nv_bfloat16* out;
  int N, H, Hkv, M, window_left, has_q2, n_qpairs;
  float scale_log2;
  int64_t o_st, o_sh;
  // paged keys (prefix-cache / chunked prefill / multi-token decode): logical key position p of
  // request n lives in page block_tables[n][p / block_size], row p % block_size
  const int32_t* block_tables;
  int64_t bt_stride;
  int paged, block_size, box_rows;
};

struct QTile { int on, rows, pmin, pmax, lo, hi; };

#ifdef HYMETA_TRACE
__device__ long long* g_swa_trace = nullptr;   // [steps][16] clock64 stamps of one CTA; tools/trace_swa_prefill.py
#define PF_TRACE(step, ev) do { if (g_swa_trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (step) < 256) \
    g_swa_trace[(step) * 16 + (ev)] = clock64(); } while (0)
#else
#define PF_TRACE(step, ev) do {} while (0)
#endif   // real key-tile range [lo, hi] (empty if hi < lo)

__global__ void __launch_bounds__(PF_THREADS, 1)
swa_prefill_kernel(const SwaPrefillParams p, const __grid_constant__ CUtensorMap tmQ,
                   const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                   const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmMK,
                   const __grid_constant__ CUtensorMap tmMV) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler
  const int h = blockIdx.x, hkv = h / (p.H / p.Hkv);
  const int qp = p.n_qpairs - 1 - blockIdx.y;   // long pairs first
  const int n = blockIdx.z;
  const bool meta_q = n == p.N;                 // pseudo-sequence of the M meta queries

  int q_row0, Lq, Lk, k_row0;
  if (meta_q) { q_row0 = 0; Lq = p.M; Lk = 0; k_row0 = 0; }
  else {
    q_row0 = p.cu_q[n]; Lq = p.cu_q[n + 1] - q_row0;
    k_row0 = p.cu_k[n]; Lk = p.cu_k[n + 1] - k_row0;
  }
  if (qp * 2 * PF_BM >= Lq) return;

  // ---- geometry shared by every role ------------------------------------------------------------
  const int shift = Lk - Lq;                                   // queries sit at the END of the keys
  const int nm = (p.M + PF_BN - 1) / PF_BN;                    // meta key tiles come first
  QTile qt[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int row0 = qp * 2 * PF_BM + i * PF_BM;
    QTile t;
    t.on = row0 < Lq;
    t.rows = t.on ? min(PF_BM, Lq - row0) : 0;
    t.pmin = shift + row0; t.pmax = t.pmin + t.rows - 1;
    t.lo = 0; t.hi = -1;
    if (t.on && !meta_q && t.pmax >= 0 && Lk > 0) {
      t.lo = p.window_left < 0 ? 0 : max(0, t.pmin - p.window_left) / PF_BN;
      t.hi = min(t.pmax, Lk - 1) / PF_BN;
    }
    qt[i] = t;
  }
  int u_lo = 0, u_hi = -1;
  if (qt[0].hi >= qt[0].lo) { u_lo = qt[0].lo; u_hi = qt[0].hi; }
  if (qt[1].hi >= qt[1].lo) {
    if (u_hi < u_lo) { u_lo = qt[1].lo; u_hi = qt[1].hi; }
    else { u_lo = min(u_lo, qt[1].lo); u_hi = max(u_hi, qt[1].hi); }
  }
  const int n_steps = nm + (u_hi >= u_lo ? u_hi - u_lo + 1 : 0);
  // lowest key position any query of this CTA can see (rows below it / at or past Lk are masked for
  // every query: in the paged form they may hold recycled-page bytes and are cleared before PV)
  const int lo_all = p.window_left < 0 ? 0 : max(0, qt[0].pmin - p.window_left);
  int last_step[2];   // last key step each query tile takes part in (-1: none)
#pragma unroll
  for (int i = 0; i < 2; ++i)
    last_step[i] = !qt[i].on ? -1 : (qt[i].hi >= qt[i].lo ? nm + qt[i].hi - u_lo : nm - 1);
  auto active = [&](int i, int t) -> bool {
    if (!qt[i].on || t >= n_steps) return false;
    if (t < nm) return true;
    const int jt = u_lo + t - nm;
    return jt >= qt[i].lo && jt <= qt[i].hi;
  };

  const uint32_t sbase = smem_u32(smem);
  const uint32_t sQ = sbase + PF_SQ, sK = sbase + PF_SK, sV = sbase + PF_SV;
  const uint32_t bars = sbase + PF_BAR;
  const uint32_t bar_q = bars, bar_kf = bars + 8, bar_ke = bar_kf + 8 * PF_NKS, bar_vf = bar_ke + 8 * PF_NKS,
                 bar_ve = bar_vf + 8 * PF_NVS, bar_s = bar_ve + 8 * PF_NVS, bar_o = bar_s + 16, bar_p = bar_o + 16;
  // bar_p: [tile][column group] -- 2 * PF_PS barriers (<= 64 bytes), ends at offset 88 + 32 + 64 = 184 < 200
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + PF_BAR + 200);

  if (tid == 0) {
    if (sbase & 1023) asm volatile("trap;");
    mbar_init(bar_q, 1);
    for (int s = 0; s < PF_NKS; ++s) { mbar_init(bar_kf + 8 * s, 1); mbar_init(bar_ke + 8 * s, 1); }
    for (int s = 0; s < PF_NVS; ++s) { mbar_init(bar_vf + 8 * s, 1); mbar_init(bar_ve + 8 * s, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_s + 8 * i, 1); mbar_init(bar_o + 8 * i, 1); }
    for (int i = 0; i < 2 * PF_PS; ++i) mbar_init(bar_p + 8 * i, 128);
    fence_mbar_init();
  }
  if (warp == 9) tmem_alloc<512>(smem_u32(tmem_slot));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  // TMEM columns: S0 [0,128) S1 [128,256) O0 [256,384) O1 [384,512); P_i aliases S_i[0,64)

  if (warp == 8) {
    // ================================================================== TMA producer
#if PF_OPT_REGS
    setmaxnreg_dec<PF_REG_LO>();
#endif
    if (elect_one_sync()) {
      const CUtensorMap* tq = meta_q ? &tmQ2 : &tmQ;
      tma_prefetch_desc(tq); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
      const int nq = qt[1].on ? 2 : 1;
      mbar_expect_tx(bar_q, nq * PF_TILE);
      for (int i = 0; i < nq; ++i) {
        const int row = q_row0 + qp * 2 * PF_BM + i * PF_BM;
        tma_load_3d(sQ + i * PF_TILE, tq, bar_q, 0, h, row);
        tma_load_3d(sQ + i * PF_TILE + 16384, tq, bar_q, 64, h, row);
      }
      const int32_t* bt = p.paged ? p.block_tables + (int64_t)n * p.bt_stride : nullptr;
      // one K or V tile: a 3-D box of the packed tensor / meta rows, or -- paged -- box_rows-row
      // boxes of the 4-D cache map, each resolved through the block table (a box never straddles a
      // page: box_rows divides block_size and logical positions are box aligned)
      auto load_tile = [&](uint32_t dst, uint32_t bar, const CUtensorMap* m3, const CUtensorMap* m4, int t) {
        mbar_expect_tx(bar, PF_TILE);
        if (t < nm || !p.paged) {
          const int row = t < nm ? t * PF_BN : k_row0 + (u_lo + t - nm) * PF_BN;
          tma_load_3d(dst, m3, bar, 0, hkv, row);
          tma_load_3d(dst + 16384, m3, bar, 64, hkv, row);
        } else {
          const int pos0 = (u_lo + t - nm) * PF_BN, rb = p.box_rows;
          for (int sub = 0; sub < PF_BN / rb; ++sub) {
            // boxes wholly outside [lo_all, Lk) are redirected to the nearest box that holds a
            // visible position (their rows are masked and their V rows cleared by the softmax warps)
            const int pos = min(max(pos0 + sub * rb, lo_all / rb * rb), Lk - 1);
            const int blk = bt[pos / p.block_size];
            const int row = (pos % p.block_size) / rb * rb;
            const uint32_t o = sub * rb * 128;
            tma_load_4d(dst + o, m4, bar, 0, hkv, row, blk);
            tma_load_4d(dst + 16384 + o, m4, bar, 64, hkv, row, blk);
          }
        }
      };
      for (int t = 0; t < n_steps; ++t) {
        {
          const int st = t % PF_NKS;
          if (t >= PF_NKS) mbar_wait(bar_ke + 8 * st, ((t / PF_NKS) - 1) & 1);
          load_tile(sK + st * PF_TILE, bar_kf + 8 * st, t < nm ? &tmMK : &tmK, &tmK, t);
          PF_TRACE(t, 12);
        }
        {
          const int st = t % PF_NVS;
          if (t >= PF_NVS) mbar_wait(bar_ve + 8 * st, ((t / PF_NVS) - 1) & 1);
          load_tile(sV + st * PF_TILE, bar_vf + 8 * st, t < nm ? &tmMV : &tmV, &tmV, t);
          PF_TRACE(t, 13);
        }
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ================================================================== MMA issuer
#if PF_OPT_REGS
    setmaxnreg_dec<PF_REG_LO>();
#endif
    if (elect_one_sync()) {
      const uint32_t idesc_s = umma_idesc_bf16(PF_BM, PF_BN, 0, 0);
      const uint32_t idesc_o = umma_idesc_bf16(PF_BM, kD, 0, 1);
      auto issue_s = [&](int i, int t) {   // S_i = Q_i K(t)^T
        const uint32_t qb = sQ + i * PF_TILE, kb = sK + (t % PF_NKS) * PF_TILE;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t da = umma_desc_sw128(qb + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          const uint64_t db = umma_desc_sw128(kb + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          umma_bf16(tmem + i * 128, da, db, idesc_s, ks > 0);
        }
        umma_commit(bar_s + 8 * i);
      };
      auto issue_pv = [&](int i, int t, bool acc, int g) {   // O_i (+)= P_i V(t), column group g
        const uint32_t vb = sV + (t % PF_NVS) * PF_TILE;
#pragma unroll
        for (int ks = g * (8 / PF_PS); ks < (g + 1) * (8 / PF_PS); ++ks) {
          const uint64_t db = umma_desc_sw128(vb + ks * 2048, 16384, 1024);
          umma_bf16_ts(tmem + 256 + i * 128, tmem + i * 128 + ks * 8, db, idesc_o, (acc || ks > 0) ? 1u : 0u);
        }
      };
      int cnt[2] = {0, 0};
      if (n_steps > 0) {
        mbar_wait(bar_q, 0);
        mbar_wait(bar_kf, 0);
        tc_fence_after();
        for (int i = 0; i < 2; ++i) if (active(i, 0)) issue_s(i, 0);
        umma_commit(bar_ke);
      }
      for (int t = 0; t < n_steps; ++t) {
        bool v_ready = false, k_ready = false;
        for (int i = 0; i < 2; ++i) {
          if (active(i, t)) {
#pragma unroll
            for (int g = 0; g < PF_PS; ++g) {
              mbar_wait(bar_p + 8 * (i * PF_PS + g), cnt[i] & 1);
              if (g == 0) PF_TRACE(t, 8 + 2 * i);
              if (!v_ready) { mbar_wait(bar_vf + 8 * (t % PF_NVS), (t / PF_NVS) & 1); v_ready = true; }
              tc_fence_after();
              issue_pv(i, t, cnt[i] > 0, g);
            }
            ++cnt[i];
            if (t == last_step[i]) umma_commit(bar_o + 8 * i);   // last key tile of this query tile
          }
          if (active(i, t + 1)) {
            if (!k_ready) { mbar_wait(bar_kf + 8 * ((t + 1) % PF_NKS), ((t + 1) / PF_NKS) & 1); k_ready = true; }
            tc_fence_after();
            issue_s(i, t + 1);
          }
          PF_TRACE(t, 9 + 2 * i);
        }
        umma_commit(bar_ve + 8 * (t % PF_NVS));
        if (t + 1 < n_steps) umma_commit(bar_ke + 8 * ((t + 1) % PF_NKS));
      }
    }
    __syncwarp();
  } else if (warp < 8) {
    // ================================================================== softmax warps
#if PF_OPT_REGS
    setmaxnreg_inc<PF_REG_HI>();
#endif
    const int i = warp >> 2;                     // query tile
    const int r = (warp & 3) * 32 + lane;        // row inside the tile == TMEM lane
    const QTile T = qt[i];
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem + i * 128 + lane_off, tO = tmem + 256 + i * 128 + lane_off;
    const int pq = T.pmin + r;                   // key position of my query
    float m_ref = -INFINITY, l_run = 0.f;        // running reference max (scaled, log2 domain) / sum
    int cnt = 0;
    if (T.on) {
      for (int t = 0; t < n_steps; ++t) {
        if (!active(i, t)) continue;
        mbar_wait(bar_s + 8 * i, cnt & 1);
        tc_fence_after();
        if ((tid & 127) == 0) PF_TRACE(t, 4 * i);
        uint32_t sr[128];
        tmem_ld32_nowait(tS + 0, *reinterpret_cast<uint32_t(*)[32]>(&sr[0]));
        tmem_ld32_nowait(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[32]));
        tmem_ld32_nowait(tS + 64, *reinterpret_cast<uint32_t(*)[32]>(&sr[64]));
        tmem_ld32_nowait(tS + 96, *reinterpret_cast<uint32_t(*)[32]>(&sr[96]));
        tmem_wait_ld();
        if ((tid & 127) == 0) PF_TRACE(t, 4 * i + 1);
        // ---- mask (tile-uniform decision)
        bool need_mask;
        int kp0, lo, hi;   // visible key positions of this row: [lo, hi]
        if (t < nm) { kp0 = t * PF_BN; need_mask = kp0 + PF_BN > p.M; lo = 0; hi = p.M - 1; }
        else {
          kp0 = (u_lo + t - nm) * PF_BN;
          need_mask = (kp0 + PF_BN - 1 > T.pmin) || (kp0 + PF_BN > Lk) ||
                      (p.window_left >= 0 && kp0 < T.pmax - p.window_left);
          hi = min(pq, Lk - 1); lo = p.window_left < 0 ? 0 : pq - p.window_left;
        }
        if (need_mask) {
#pragma unroll
          for (int c = 0; c < 128; ++c) {
            const int kp = kp0 + c;
            if (kp < lo || kp > hi) sr[c] = 0xff800000u;   // -inf
          }
        }
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int c = 0; c < 128; c += 4) {
          mx0 = fmaxf(mx0, __uint_as_float(sr[c])); mx1 = fmaxf(mx1, __uint_as_float(sr[c + 1]));
          mx2 = fmaxf(mx2, __uint_as_float(sr[c + 2])); mx3 = fmaxf(mx3, __uint_as_float(sr[c + 3]));
        }
        const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;
        // ---- lazy rescale of O (in TMEM) when the row maximum jumps
        const bool grow = mx > m_ref + PF_RESCALE_T || (m_ref == -INFINITY && mx > -INFINITY);
        if (__any_sync(0xffffffffu, grow)) {
          float alpha = 1.f;
          if (grow) {
            alpha = (m_ref == -INFINITY) ? 1.f : exp2f(m_ref - mx);
            m_ref = mx;
            l_run *= alpha;
          }
          if (cnt > 0) {   // O_i holds PV(0..cnt-1): all complete, since S_i(t) completed after them
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint32_t o[32];
              tmem_ld32_nowait(tO + 32 * c, o);
              tmem_wait_ld();
#pragma unroll
              for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * alpha);
              tmem_st32(tO + 32 * c, o);
            }
          }
        }
        if ((tid & 127) == 0) PF_TRACE(t, 4 * i + 2);
        const float msub = (m_ref == -INFINITY) ? 0.f : m_ref;
        if (p.paged && t >= nm && (kp0 + PF_BN > Lk || kp0 < lo_all)) {
          // paged boundary tile: key rows outside [lo_all, Lk) are masked for every query (P = 0)
          // but may hold recycled-page bytes (NaN patterns), and 0 * NaN would poison O -- clear
          // those V rows in the landed stage before PV is issued (generic -> async proxy fence)
          const int st = t % PF_NVS;
          mbar_wait(bar_vf + 8 * st, (t / PF_NVS) & 1);
          const int kp = kp0 + r;
          if (kp >= Lk || kp < lo_all) {
            uint8_t* row = smem + PF_SV + st * PF_TILE + r * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              *reinterpret_cast<uint4*>(row + c * 16) = make_uint4(0, 0, 0, 0);
              *reinterpret_cast<uint4*>(row + 16384 + c * 16) = make_uint4(0, 0, 0, 0);
            }
          }
          fence_proxy_async_smem();
        }
        float rs0 = 0.f, rs1 = 0.f, rs2 = 0.f, rs3 = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t w[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float a = exp2f(fmaf(__uint_as_float(sr[32 * c + 2 * j]), p.scale_log2, -msub));
            const float b = exp2f(fmaf(__uint_as_float(sr[32 * c + 2 * j + 1]), p.scale_log2, -msub));
            if (j & 1) { rs2 += a; rs3 += b; } else { rs0 += a; rs1 += b; }
            w[j] = pack_bf16(a, b);
          }
          tmem_st16(tS + 16 * c, w);
          if ((c + 1) % (4 / PF_PS) == 0) {   // column group complete: hand it to the MMA warp
            tmem_wait_st();
            tc_fence_before();
            mbar_arrive(bar_p + 8 * (i * PF_PS + c / (4 / PF_PS)));
          }
        }
        l_run += (rs0 + rs1) + (rs2 + rs3);
        if ((tid & 127) == 0) PF_TRACE(t, 4 * i + 3);
        ++cnt;
      }
      // ---- epilogue: normalise, bf16, one 256-byte row per thread
      if (cnt > 0) { mbar_wait(bar_o + 8 * i, 0); tc_fence_after(); }
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
      const int64_t orow = (meta_q ? 0 : (int64_t)(p.has_q2 ? p.M : 0) + q_row0) + qp * 2 * PF_BM + i * PF_BM + r;
      __nv_bfloat16* dst = p.out + orow * p.o_st + (int64_t)h * p.o_sh;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t o[32];
        if (cnt > 0) { tmem_ld32_nowait(tO + 32 * c, o); tmem_wait_ld(); }
        else {
#pragma unroll
          for (int j = 0; j < 32; ++j) o[j] = 0u;
        }
        if (r < T.rows) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 v;
            v.x = pack_bf16(__uint_as_float(o[8 * j + 0]) * inv, __uint_as_float(o[8 * j + 1]) * inv);
            v.y = pack_bf16(__uint_as_float(o[8 * j + 2]) * inv, __uint_as_float(o[8 * j + 3]) * inv);
            v.z = pack_bf16(__uint_as_float(o[8 * j + 4]) * inv, __uint_as_float(o[8 * j + 5]) * inv);
            v.w = pack_bf16(__uint_as_float(o[8 * j + 6]) * inv, __uint_as_float(o[8 * j + 7]) * inv);
            stg_v4(dst + 32 * c + 8 * j, v);
          }
        }
      }
    }
  }
#if PF_OPT_REGS
  else setmaxnreg_dec<PF_REG_LO>();   // warps 10-11 only complete the third warpgroup
#endif
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<512>(tmem);
}

}  // namespace hymeta

using namespace hymeta;

#ifdef HYMETA_TRACE
extern "C" int hymeta_debug_set_swa_trace(void* buf) {
  return (int)cudaMemcpyToSymbol(g_swa_trace, &buf, sizeof(buf));
}
#endif

extern "C" int hymeta_swa_meta_prefill_varlen(
    const void* q1, const void* k1, const void* v1, const void* q2, const void* meta_k,
    const void* meta_v, void* out, const int32_t* cu_q, const int32_t* cu_k,
    const int32_t* block_tables, int64_t bt_stride, int block_size, int N, int Tq, int Tk, int H,
    int Hkv, int D, int M, int max_seqlen_q, int window_left, float scale, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t q2_st, int64_t q2_sh,
    int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, int64_t o_st, int64_t o_sh,
    int64_t kc_sblk, int64_t vc_sblk, void* stream) {
  if (N < 0 || Tq < 0 || Tk < 0 || H <= 0 || Hkv <= 0 || M < 0 || !out) return HYMETA_ERR_INVALID_ARG;
  if (D != kD || H % Hkv) return HYMETA_ERR_UNSUPPORTED;
  const bool paged = block_tables != nullptr;
  if (paged && (block_size < 16 || block_size % 16 || Tk <= 0 || Tk % block_size || bt_stride < 0))
    return block_size > 0 && block_size % 16 ? HYMETA_ERR_UNSUPPORTED : HYMETA_ERR_INVALID_ARG;
  if (M > 0 && (!meta_k || !meta_v)) return HYMETA_ERR_INVALID_ARG;
  if (q2 && M == 0) return HYMETA_ERR_INVALID_ARG;
  const bool real = N > 0 && Tq > 0;
  if (real && (!q1 || !k1 || !v1 || !cu_q || !cu_k)) return HYMETA_ERR_INVALID_ARG;
  if (!real && !q2) return HYMETA_OK;
  if (N > 65534 || H > 65535) return HYMETA_ERR_UNSUPPORTED;

  CUtensorMap tmQ, tmK, tmV, tmQ2, tmMK, tmMV;
  memset(&tmQ, 0, sizeof(tmQ)); tmK = tmV = tmQ2 = tmMK = tmMV = tmQ;
  auto mk3 = [&](CUtensorMap* tm, const void* base, int heads, int rows, int64_t st, int64_t sh,
                 uint32_t box_rows) {
    const uint64_t dims[3] = {(uint64_t)kD, (uint64_t)heads, (uint64_t)rows};
    const int64_t strides[3] = {1, sh, st};
    const uint32_t box[3] = {64, 1, box_rows};
    return make_tmap_bf16(tm, base, 3, dims, strides, box, true);
  };
  // paged: [D, Hkv, block_size, num_blocks] over the caches, boxes of box_rows key rows
  const int box_rows = !paged ? PF_BN : block_size % 128 == 0 ? 128 : block_size % 64 == 0 ? 64
                                        : block_size % 32 == 0 ? 32 : 16;
  auto mk4 = [&](CUtensorMap* tm, const void* base, int64_t s_h, int64_t s_tok, int64_t s_blk) {
    const uint64_t dims[4] = {(uint64_t)kD, (uint64_t)Hkv, (uint64_t)block_size, (uint64_t)(Tk / block_size)};
    const int64_t str[4] = {1, s_h, s_tok, s_blk};
    const uint32_t box[4] = {64, 1, (uint32_t)box_rows, 1};
    return make_tmap_bf16(tm, base, 4, dims, str, box, true);
  };
  int rc = HYMETA_OK;
  if (real) {
    if ((rc = mk3(&tmQ, q1, H, Tq, q_st, q_sh, PF_BM))) return rc;
    if (paged) {
      if ((rc = mk4(&tmK, k1, k_sh, k_st, kc_sblk))) return rc;
      if ((rc = mk4(&tmV, v1, v_sh, v_st, vc_sblk))) return rc;
    } else {
      if ((rc = mk3(&tmK, k1, Hkv, Tk, k_st, k_sh, PF_BN))) return rc;
      if ((rc = mk3(&tmV, v1, Hkv, Tk, v_st, v_sh, PF_BN))) return rc;
    }
  }
  if (M > 0) {
    if ((rc = mk3(&tmMK, meta_k, Hkv, M, mk_st, mk_sh, PF_BN))) return rc;
    if ((rc = mk3(&tmMV, meta_v, Hkv, M, mv_st, mv_sh, PF_BN))) return rc;
    if (q2 && (rc = mk3(&tmQ2, q2, H, M, q2_st, q2_sh, PF_BM))) return rc;
  }
  if (!real) { tmQ = tmQ2; tmK = tmMK; tmV = tmMV; }   // never dereferenced, but must be valid maps
  if (!q2) tmQ2 = tmQ;
  if (M == 0) { tmMK = tmQ; tmMV = tmQ; }               // never dereferenced (no meta tiles)

  SwaPrefillParams p;
  p.cu_q = cu_q; p.cu_k = cu_k; p.out = (__nv_bfloat16*)out;
  p.N = real ? N : 0; p.H = H; p.Hkv = Hkv; p.M = M; p.window_left = window_left;
  p.has_q2 = q2 ? 1 : 0;
  int msq = real ? max_seqlen_q : 0;
  if (q2 && M > msq) msq = M;
  if (msq <= 0) msq = real ? Tq : M;
  p.n_qpairs = (msq + 2 * PF_BM - 1) / (2 * PF_BM);
  p.scale_log2 = scale * 1.4426950408889634f;
  p.o_st = o_st; p.o_sh = o_sh;
  p.block_tables = block_tables; p.bt_stride = bt_stride;
  p.paged = paged && real ? 1 : 0; p.block_size = paged ? block_size : 1; p.box_rows = box_rows;

  static DynSmemAttr attr;
  attr.ensure(swa_prefill_kernel, PF_SMEM);
  dim3 grid(H, p.n_qpairs, p.N + (q2 ? 1 : 0));
  swa_prefill_kernel<<<grid, PF_THREADS, PF_SMEM, (cudaStream_t)stream>>>(p, tmQ, tmK, tmV, tmQ2, tmMK, tmMV);
  count_launch();
  return check_launch();
}
