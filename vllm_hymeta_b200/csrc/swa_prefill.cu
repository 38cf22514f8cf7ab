// SWA + meta-prefix varlen prefill on tcgen05 / TMEM / TMA (a6/a7).
//
// One launch replaces the reference's two flash-attn launches (windowed causal pass + non-causal
// meta pass), the concat of q, the 6-kernel TorchScript LSE merge and the output concat
// (fa_kernel_metatoken.py:198-285): the meta keys are simply the first key tiles of every query
// tile's single online softmax.
//
// PERSISTENT: one CTA per SM (all 512 TMEM columns, 224 KB of smem) walks work items
// item = (q-head, pair of 128-row query tiles, request), request-major with each request's longest
// pairs first, CTA c taking items c, c + grid, ... -- consecutive items are the heads of one (pair,
// request), so the G heads of a GQA group run side by side on neighbouring SMs and a request's K/V
// stay in L2 while its pairs are walked.  Barriers, the K/V
// rings and TMEM live across items, and the roles run decoupled, so the next item's Q/K loads and
// its first S = Q K^T are in flight while the softmax warps still normalise and store the current
// item's output: the per-item fill/drain (the cost that dominated short prompts) is overlapped.
//
// Warp-specialised, 3 warpgroups, setmaxnreg gives the softmax warps 208 registers (their 128
// scores stay in registers; the hot loop has no spills):
//   warp 8      TMA producer: Q0/Q1 per item, then K tiles (2-stage ring) and V tiles (3-stage
//               ring), SWIZZLE_128B boxes straight from the packed [T, Hkv, D] tensors or -- paged --
//               from the [D, Hkv, block_size, num_blocks] cache through the block table
//   warp 9      MMA issuer (one lane):  S_i = Q_i K^T          tcgen05.mma SS, fp32 S in TMEM
//                                       O_i += P_i V           tcgen05.mma TS: A = bf16 P in TMEM
//                                                              (aliases S_i), B = V MN-major as landed
//   warps 0-3   softmax of query tile 0, warps 4-7 of tile 1 (thread = query row = TMEM lane):
//               one tcgen05.ld of the 128 scores, masks only on boundary tiles, exp2 in the log2
//               domain, bf16 P stored back over S with tcgen05.st and handed to the MMA warp in
//               PF_PS column groups (PV of the first group runs under the exp2 of the second).  O
//               stays in TMEM for the whole key loop; it is rescaled in place only when a row
//               maximum grows by more than 2^8 (lazy rescale).
//   warps 10-11 idle (they complete the third warpgroup for setmaxnreg)
// The issue order  PV0(t) S0(t+1) PV1(t) S1(t+1)  lets tile 1's softmax run under tile 0's MMAs and
// vice versa; at an item's last step S_i(t+1) is the NEXT item's first S.  MMAs of one CTA complete
// in order, so "S_i(t) has landed" also means O_i holds every PV up to t-1 and the rescale needs no
// extra barrier; O_i is handed back to the MMA warp (bar_of) once the epilogue has read it.
#include "umma.cuh"

namespace hymeta {

constexpr int PF_BM = 128;    // query rows per tile (two tiles per item)
constexpr int PF_BN = 128;    // keys per tile
constexpr int PF_NKS = 2;     // K ring stages
#ifndef PF_OPT_EPI
#define PF_OPT_EPI 1      // 1: output rows staged per warp through smem and written as full 128-byte lines; 0: one 256-byte row per thread
#endif
#ifndef PF_OPT_MASK
#define PF_OPT_MASK 1     // 1: boundary tiles masked from a per-row 128-bit visibility mask; 0: two compares per score
#endif
#ifndef PF_NVS_
#define PF_NVS_ (PF_OPT_EPI ? 2 : 3)
#endif
constexpr int PF_NVS = PF_NVS_;   // V ring stages (the third stage's 32 KB is the output staging area when PF_OPT_EPI)
constexpr int PF_TILE = 32768;                          // one [128][128] bf16 tile: 2 halves x [128][64]
constexpr int PF_SQ = 0;                                // Q0, Q1
constexpr int PF_SK = 2 * PF_TILE;
constexpr int PF_SV = PF_SK + PF_NKS * PF_TILE;
constexpr int PF_SO = PF_SV + PF_NVS * PF_TILE;         // output staging: 8 warps x 4 KB
constexpr int PF_BAR = PF_SO + (PF_OPT_EPI ? 8 * 4096 : 0);   // mbarriers + tmem ptr
constexpr int PF_SMEM = PF_BAR + 256;
#ifndef PF_OPT_PSPLIT
#define PF_OPT_PSPLIT 2   // P handed to the MMA warp in this many column groups (1, 2 or 4), each with its own mbarrier
#endif
#ifndef PF_REG_LO
#define PF_REG_LO 88      // producer / MMA / idle warps
#define PF_REG_HI 208     // softmax warps: 256 * 208 + 128 * 88 = 64512 = 384 * 168 (the launch allocation)
#endif
constexpr int PF_THREADS = 384;
constexpr int PF_PS = PF_OPT_PSPLIT;
constexpr float PF_RESCALE_T = 8.f;                     // lazy rescale threshold (log2 domain)

struct SwaPrefillParams {
  const int32_t *cu_q, *cu_k;
  __nv_bfloat16* out;
  int N, H, Hkv, M, window_left, has_q2, n_qpairs;
  int n_z, n_items;             // requests incl. the meta-query pseudo-request; H * n_qpairs * n_z
  float scale_log2;
  int64_t o_st, o_sh;
  // paged keys (prefix-cache / chunked prefill / multi-token decode): logical key position p of
  // request n lives in page block_tables[n][p / block_size], row p % block_size
  const int32_t* block_tables;
  int64_t bt_stride;
  int paged, block_size, box_rows;
  // f1: per (output row, head) sum of squares of the bf16 output row, [rows, H] fp32, or null --
  // the branch merge's RMS norm partials, emitted here so that the merge need not read x for them
  float* sumsq;
};

struct QTile { int on, rows, pmin, pmax, lo, hi; };   // real key-tile range [lo, hi] (empty if hi < lo)

// Geometry of one work item, recomputed by every role from the same inputs.
struct ItemGeo {
  int valid;          // the item exists and holds query rows
  int n_steps;        // key steps (meta tiles first, then real tiles u_lo ..); 0: no visible key at all
  int h, hkv, n, meta_q;
  int q_row;          // first query row of the pair in the packed q tensor (meta_q: in q2)
  int64_t o_row;      // first output row of the pair
  int k_row0, Lk, nm, u_lo, lo_all;
  QTile qt[2];
  int first_step[2], last_step[2];   // first / last key step each query tile takes part in (-1: none)
};

__device__ __forceinline__ ItemGeo make_geo(const SwaPrefillParams& p, int idx) {
  ItemGeo g;
  g.valid = 0; g.n_steps = 0; g.h = g.hkv = g.n = g.meta_q = g.q_row = 0; g.o_row = 0;
  g.k_row0 = g.Lk = g.nm = g.u_lo = g.lo_all = 0;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    g.qt[i].on = g.qt[i].rows = g.qt[i].pmin = g.qt[i].lo = 0; g.qt[i].pmax = g.qt[i].hi = -1;
    g.first_step[i] = g.last_step[i] = -1;
  }
  if (idx >= p.n_items) return g;
  const int r = idx / p.H;
  g.h = idx - r * p.H; g.hkv = g.h / (p.H / p.Hkv);
  // request-major, inside a request the long (late) pairs first: the ~5 (pair, request) groups
  // in flight at any time then belong to one or two requests, whose K/V (16 MB at 8K tokens) stay
  // in L2 across all of the request's pairs.  (Pair-level-major order re-read every K/V window
  // from DRAM once per level: 20 GB instead of 4.8 GB at config 2.)
  g.n = r / p.n_qpairs;
  const int qp = p.n_qpairs - 1 - (r - g.n * p.n_qpairs);
  g.meta_q = g.n == p.N;                        // pseudo-request of the M meta queries
  int q_row0, Lq;
  if (g.meta_q) { q_row0 = 0; Lq = p.M; g.Lk = 0; g.k_row0 = 0; }
  else {
    q_row0 = __ldg(p.cu_q + g.n); Lq = __ldg(p.cu_q + g.n + 1) - q_row0;
    g.k_row0 = __ldg(p.cu_k + g.n); g.Lk = __ldg(p.cu_k + g.n + 1) - g.k_row0;
  }
  if (qp * 2 * PF_BM >= Lq) return g;
  g.valid = 1;
  g.q_row = q_row0 + qp * 2 * PF_BM;
  g.o_row = (g.meta_q ? 0 : (int64_t)(p.has_q2 ? p.M : 0) + q_row0) + qp * 2 * PF_BM;
  const int shift = g.Lk - Lq;                                 // queries sit at the END of the keys
  g.nm = (p.M + PF_BN - 1) / PF_BN;                            // meta key tiles come first
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int row0 = qp * 2 * PF_BM + i * PF_BM;
    QTile t;
    t.on = row0 < Lq;
    t.rows = t.on ? min(PF_BM, Lq - row0) : 0;
    t.pmin = shift + row0; t.pmax = t.pmin + t.rows - 1;
    t.lo = 0; t.hi = -1;
    if (t.on && !g.meta_q && t.pmax >= 0 && g.Lk > 0) {
      t.lo = p.window_left < 0 ? 0 : max(0, t.pmin - p.window_left) / PF_BN;
      t.hi = min(t.pmax, g.Lk - 1) / PF_BN;
    }
    g.qt[i] = t;
  }
  int u_lo = 0, u_hi = -1;
  if (g.qt[0].hi >= g.qt[0].lo) { u_lo = g.qt[0].lo; u_hi = g.qt[0].hi; }
  if (g.qt[1].hi >= g.qt[1].lo) {
    if (u_hi < u_lo) { u_lo = g.qt[1].lo; u_hi = g.qt[1].hi; }
    else { u_lo = min(u_lo, g.qt[1].lo); u_hi = max(u_hi, g.qt[1].hi); }
  }
  g.u_lo = u_lo;
  g.n_steps = g.nm + (u_hi >= u_lo ? u_hi - u_lo + 1 : 0);
  // lowest key position any query of this item can see (rows below it / at or past Lk are masked for
  // every query: in the paged form they may hold recycled-page bytes and are cleared before PV)
  g.lo_all = p.window_left < 0 ? 0 : max(0, g.qt[0].pmin - p.window_left);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const bool real = g.qt[i].hi >= g.qt[i].lo;
    g.last_step[i] = !g.qt[i].on ? -1 : (real ? g.nm + g.qt[i].hi - u_lo : g.nm - 1);
    g.first_step[i] = !g.qt[i].on ? -1 : (g.nm > 0 ? 0 : (real ? g.qt[i].lo - u_lo : -1));
  }
  return g;
}

// does query tile i take part in key step t of the item
__device__ __forceinline__ bool geo_active(const ItemGeo& g, const QTile& q, int t) {
  if (!q.on || t >= g.n_steps) return false;
  if (t < g.nm) return true;
  const int jt = g.u_lo + t - g.nm;
  return jt >= q.lo && jt <= q.hi;
}
__device__ __forceinline__ bool geo_active(const ItemGeo& g, int i, int t) {   // i must be a compile-time constant
  return geo_active(g, g.qt[i], t);
}

#ifdef HYMETA_TRACE
__device__ long long* g_swa_trace = nullptr;   // [steps][16] clock64 stamps of CTA 0's first 256 key steps; tools/trace_swa_prefill.py
#define PF_TRACE(step, ev) do { if (g_swa_trace && blockIdx.x == 0 && (step) < 256) \
    g_swa_trace[(step) * 16 + (ev)] = clock64(); } while (0)
#else
#define PF_TRACE(step, ev) do {} while (0)
#endif

__global__ void __launch_bounds__(PF_THREADS, 1)
swa_prefill_kernel(const SwaPrefillParams p, const __grid_constant__ CUtensorMap tmQ,
                   const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                   const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmMK,
                   const __grid_constant__ CUtensorMap tmMV) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);   // warp-uniform for the compiler
  const int G = gridDim.x;

  const uint32_t sbase = smem_u32(smem);
  const uint32_t sQ = sbase + PF_SQ, sK = sbase + PF_SK, sV = sbase + PF_SV;
  const uint32_t bars = sbase + PF_BAR;
  // mbarriers (8 bytes each): Q full/empty per tile, K/V ring full/empty, S landed per tile, PV of
  // the item complete per tile (bar_o), O read back by the epilogue per tile (bar_of), P column
  // groups [tile][group]; 2+2+2+2+3+3+2+2+2+8 = 28 slots = 224 bytes, tmem pointer at +248
  const uint32_t bar_qf = bars, bar_qe = bar_qf + 16, bar_kf = bar_qe + 16, bar_ke = bar_kf + 8 * PF_NKS,
                 bar_vf = bar_ke + 8 * PF_NKS, bar_ve = bar_vf + 8 * PF_NVS, bar_s = bar_ve + 8 * PF_NVS,
                 bar_o = bar_s + 16, bar_of = bar_o + 16, bar_p = bar_of + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + PF_BAR + 248);

  if (tid == 0) {
    if (sbase & 1023) asm volatile("trap;");
    for (int i = 0; i < 2; ++i) { mbar_init(bar_qf + 8 * i, 1); mbar_init(bar_qe + 8 * i, 1); }
    for (int s = 0; s < PF_NKS; ++s) { mbar_init(bar_kf + 8 * s, 1); mbar_init(bar_ke + 8 * s, 1); }
    for (int s = 0; s < PF_NVS; ++s) { mbar_init(bar_vf + 8 * s, 1); mbar_init(bar_ve + 8 * s, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_s + 8 * i, 1); mbar_init(bar_o + 8 * i, 1); mbar_init(bar_of + 8 * i, 128); }
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
    setmaxnreg_dec<PF_REG_LO>();
    if (elect_one_sync()) {
      tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
      int gt = 0;               // key steps issued so far (ring position of the item's first K / V tile)
      int qc[2] = {0, 0};       // Q loads issued per tile
      for (int it = blockIdx.x; it < p.n_items; it += G) {
        const ItemGeo g = make_geo(p, it);
        if (!g.valid || g.n_steps == 0) continue;
        const CUtensorMap* tq = g.meta_q ? &tmQ2 : &tmQ;
        // Q_i goes out right before the K tile of the first step that needs it (tile 1 of the
        // previous item is still using its Q while this item's Q_0 / K / V are already in flight)
        auto load_q = [&](int i) {
          if (qc[i] > 0) mbar_wait(bar_qe + 8 * i, (qc[i] - 1) & 1);   // last S of the previous use has read it
          mbar_expect_tx(bar_qf + 8 * i, PF_TILE);
          const int row = g.q_row + i * PF_BM;
          tma_load_3d(sQ + i * PF_TILE, tq, bar_qf + 8 * i, 0, g.h, row);
          tma_load_3d(sQ + i * PF_TILE + 16384, tq, bar_qf + 8 * i, 64, g.h, row);
          ++qc[i];
        };
        const int32_t* bt = p.paged ? p.block_tables + (int64_t)g.n * p.bt_stride : nullptr;
        // one K or V tile: a 3-D box of the packed tensor / meta rows, or -- paged -- box_rows-row
        // boxes of the 4-D cache map, each resolved through the block table (a box never straddles a
        // page: box_rows divides block_size and logical positions are box aligned)
        auto load_tile = [&](uint32_t dst, uint32_t bar, const CUtensorMap* m3, const CUtensorMap* m4, int t) {
          mbar_expect_tx(bar, PF_TILE);
          if (t < g.nm || !p.paged) {
            const int row = t < g.nm ? t * PF_BN : g.k_row0 + (g.u_lo + t - g.nm) * PF_BN;
            tma_load_3d(dst, m3, bar, 0, g.hkv, row);
            tma_load_3d(dst + 16384, m3, bar, 64, g.hkv, row);
          } else {
            const int pos0 = (g.u_lo + t - g.nm) * PF_BN, rb = p.box_rows;
            for (int sub = 0; sub < PF_BN / rb; ++sub) {
              // boxes wholly outside [lo_all, Lk) are redirected to the nearest box that holds a
              // visible position (their rows are masked and their V rows cleared by the softmax warps)
              const int pos = min(max(pos0 + sub * rb, g.lo_all / rb * rb), g.Lk - 1);
              const int blk = bt[pos / p.block_size];
              const int row = (pos % p.block_size) / rb * rb;
              const uint32_t o = sub * rb * 128;
              tma_load_4d(dst + o, m4, bar, 0, g.hkv, row, blk);
              tma_load_4d(dst + 16384 + o, m4, bar, 64, g.hkv, row, blk);
            }
          }
        };
        for (int t = 0; t < g.n_steps; ++t) {
          const int gi = gt + t;
          if (t == g.first_step[0]) load_q(0);
          if (t == g.first_step[1]) load_q(1);
          {
            const int st = gi % PF_NKS;
            if (gi >= PF_NKS) mbar_wait(bar_ke + 8 * st, ((gi / PF_NKS) - 1) & 1);
            load_tile(sK + st * PF_TILE, bar_kf + 8 * st, t < g.nm ? &tmMK : &tmK, &tmK, t);
            PF_TRACE(gi, 12);
          }
          {
            const int st = gi % PF_NVS;
            if (gi >= PF_NVS) mbar_wait(bar_ve + 8 * st, ((gi / PF_NVS) - 1) & 1);
            load_tile(sV + st * PF_TILE, bar_vf + 8 * st, t < g.nm ? &tmMV : &tmV, &tmV, t);
            PF_TRACE(gi, 13);
          }
        }
        gt += g.n_steps;
      }
    }
    __syncwarp();
  } else if (warp == 9) {
    // ================================================================== MMA issuer
    setmaxnreg_dec<PF_REG_LO>();
    if (elect_one_sync()) {
      const uint32_t idesc_s = umma_idesc_bf16(PF_BM, PF_BN, 0, 0);
      const uint32_t idesc_o = umma_idesc_bf16(PF_BM, kD, 0, 1);
      int cnt[2] = {0, 0};      // softmax hand-offs consumed per tile (parity of bar_p)
      int qc[2] = {0, 0};       // Q tiles consumed per tile (parity of bar_qf)
      int oc[2] = {0, 0};       // items whose O_i was produced (parity of bar_of)
      // S_i = Q_i K^T for key step t of item g, whose first key step is ring position gt0
      auto issue_s = [&](const ItemGeo& g, int i, int t, int gt0) {
        if (t == g.first_step[i]) { mbar_wait(bar_qf + 8 * i, qc[i] & 1); ++qc[i]; }
        tc_fence_after();
        const uint32_t qb = sQ + i * PF_TILE, kb = sK + ((gt0 + t) % PF_NKS) * PF_TILE;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t da = umma_desc_sw128(qb + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          const uint64_t db = umma_desc_sw128(kb + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          umma_bf16(tmem + i * 128, da, db, idesc_s, ks > 0);
        }
        umma_commit(bar_s + 8 * i);
        if (t == g.last_step[i]) umma_commit(bar_qe + 8 * i);   // Q_i is free once this S has completed
      };
      auto issue_pv = [&](int i, int gi, bool acc, int grp) {   // O_i (+)= P_i V, column group grp
        const uint32_t vb = sV + (gi % PF_NVS) * PF_TILE;
#pragma unroll
        for (int ks = grp * (8 / PF_PS); ks < (grp + 1) * (8 / PF_PS); ++ks) {
          const uint64_t db = umma_desc_sw128(vb + ks * 2048, 16384, 1024);
          umma_bf16_ts(tmem + 256 + i * 128, tmem + i * 128 + ks * 8, db, idesc_o, (acc || ks > 0) ? 1u : 0u);
        }
      };
      auto skip = [&](const ItemGeo& g) { return !g.valid || g.n_steps == 0; };
      int it = blockIdx.x;
      ItemGeo g = make_geo(p, it);
      while (it < p.n_items && skip(g)) { it += G; g = make_geo(p, it); }
      bool have = it < p.n_items;
      int gt = 0;
      if (have) {   // first item: its step-0 S has no previous item to ride on
        mbar_wait(bar_kf, 0);
#pragma unroll
        for (int i = 0; i < 2; ++i) if (geo_active(g, i, 0)) issue_s(g, i, 0, 0);
        umma_commit(bar_ke);
      }
      while (have) {
        int nit = it + G;
        ItemGeo ng = make_geo(p, nit);
        while (nit < p.n_items && skip(ng)) { nit += G; ng = make_geo(p, nit); }
        const bool have_next = nit < p.n_items;
        int ci[2] = {0, 0};   // PVs issued for this item
        for (int t = 0; t < g.n_steps; ++t) {
          const int gi = gt + t;
          const bool last = t == g.n_steps - 1;
          bool v_ready = false, k_ready = false;
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            if (geo_active(g, i, t)) {
              // O_i of the previous item has been read back by the epilogue
              if (ci[i] == 0 && oc[i] > 0) mbar_wait(bar_of + 8 * i, (oc[i] - 1) & 1);
#pragma unroll
              for (int grp = 0; grp < PF_PS; ++grp) {
                mbar_wait(bar_p + 8 * (i * PF_PS + grp), cnt[i] & 1);
                if (grp == 0) PF_TRACE(gi, 8 + 2 * i);
                if (!v_ready) { mbar_wait(bar_vf + 8 * (gi % PF_NVS), (gi / PF_NVS) & 1); v_ready = true; }
                tc_fence_after();
                issue_pv(i, gi, ci[i] > 0, grp);
              }
              ++cnt[i]; ++ci[i];
              if (t == g.last_step[i]) { umma_commit(bar_o + 8 * i); ++oc[i]; }   // last key tile of this query tile
            }
            // the next S of this tile: step t+1 of this item, or step 0 of the next item
            const bool s_here = !last && geo_active(g, i, t + 1);
            const bool s_next = last && have_next && geo_active(ng, i, 0);
            if (s_here || s_next) {
              if (!k_ready) { mbar_wait(bar_kf + 8 * ((gi + 1) % PF_NKS), ((gi + 1) / PF_NKS) & 1); k_ready = true; }
              if (s_here) issue_s(g, i, t + 1, gt); else issue_s(ng, i, 0, gt + g.n_steps);
            }
            PF_TRACE(gi, 9 + 2 * i);
          }
          umma_commit(bar_ve + 8 * (gi % PF_NVS));
          if (!last || have_next) umma_commit(bar_ke + 8 * ((gi + 1) % PF_NKS));
        }
        gt += g.n_steps;
        it = nit; g = ng; have = have_next;
      }
    }
    __syncwarp();
  } else if (warp < 8) {
    // ================================================================== softmax warps
    setmaxnreg_inc<PF_REG_HI>();
    const int i = warp >> 2;                     // query tile
    const int r = (warp & 3) * 32 + lane;        // row inside the tile == TMEM lane
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem + i * 128 + lane_off, tO = tmem + 256 + i * 128 + lane_off;
    int cnt = 0;      // S tiles consumed (parity of bar_s / bar_p)
    int oc = 0;       // items with an O (parity of bar_o)
    int gt = 0;       // ring position of the item's first key step
    for (int it = blockIdx.x; it < p.n_items; it += G) {
      const ItemGeo g = make_geo(p, it);
      if (!g.valid) continue;
      const QTile T = i ? g.qt[1] : g.qt[0];
      if (!T.on) { gt += g.n_steps; continue; }
      const int nm = g.nm, Lk = g.Lk;
      const int pq = T.pmin + r;                   // key position of my query
      float m_ref = -INFINITY, l_run = 0.f;        // running reference max (scaled, log2 domain) / sum
      int ci = 0;                                  // key steps of this item done
      for (int t = 0; t < g.n_steps; ++t) {
        if (!geo_active(g, T, t)) continue;
        const int gi = gt + t;
        mbar_wait(bar_s + 8 * i, cnt & 1);
        tc_fence_after();
        if ((tid & 127) == 0) PF_TRACE(gi, 4 * i);
        uint32_t sr[128];
        tmem_ld32_nowait(tS + 0, *reinterpret_cast<uint32_t(*)[32]>(&sr[0]));
        tmem_ld32_nowait(tS + 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[32]));
        tmem_ld32_nowait(tS + 64, *reinterpret_cast<uint32_t(*)[32]>(&sr[64]));
        tmem_ld32_nowait(tS + 96, *reinterpret_cast<uint32_t(*)[32]>(&sr[96]));
        tmem_wait_ld();
        if ((tid & 127) == 0) PF_TRACE(gi, 4 * i + 1);
        // ---- mask (tile-uniform decision)
        bool need_mask;
        int kp0, lo, hi;   // visible key positions of this row: [lo, hi]
        if (t < nm) { kp0 = t * PF_BN; need_mask = kp0 + PF_BN > p.M; lo = 0; hi = p.M - 1; }
        else {
          kp0 = (g.u_lo + t - nm) * PF_BN;
          need_mask = (kp0 + PF_BN - 1 > T.pmin) || (kp0 + PF_BN > Lk) ||
                      (p.window_left >= 0 && kp0 < T.pmax - p.window_left);
          hi = min(pq, Lk - 1); lo = p.window_left < 0 ? 0 : pq - p.window_left;
        }
        if (need_mask) {
#if PF_OPT_MASK
          // visible columns of my row are [lo - kp0, hi - kp0]: one 32-bit visibility word per 32
          // columns, tested bit by bit (one predicate + select per score instead of two compares)
          const int c_lo = lo - kp0, c_hi = hi - kp0;
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            const int a = min(max(c_lo - 32 * w, 0), 32), b = min(max(c_hi + 1 - 32 * w, 0), 32);   // [a, b)
            const uint32_t below_b = b >= 32 ? 0xffffffffu : (1u << b) - 1u;
            const uint32_t from_a = a >= 32 ? 0u : 0xffffffffu << a;
            const uint32_t vis = below_b & from_a;
#pragma unroll
            for (int c = 0; c < 32; ++c)
              if (!(vis & (1u << c))) sr[32 * w + c] = 0xff800000u;   // -inf
          }
#else
#pragma unroll
          for (int c = 0; c < 128; ++c) {
            const int kp = kp0 + c;
            if (kp < lo || kp > hi) sr[c] = 0xff800000u;   // -inf
          }
#endif
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
          if (ci > 0) {   // O_i holds PV(0..ci-1): all complete, since S_i(t) completed after them
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
        if ((tid & 127) == 0) PF_TRACE(gi, 4 * i + 2);
        const float msub = (m_ref == -INFINITY) ? 0.f : m_ref;
        if (p.paged && t >= nm && (kp0 + PF_BN > Lk || kp0 < g.lo_all)) {
          // paged boundary tile: key rows outside [lo_all, Lk) are masked for every query (P = 0)
          // but may hold recycled-page bytes (NaN patterns), and 0 * NaN would poison O -- clear
          // those V rows in the landed stage before PV is issued (generic -> async proxy fence)
          const int st = gi % PF_NVS;
          mbar_wait(bar_vf + 8 * st, (gi / PF_NVS) & 1);
          const int kp = kp0 + r;
          if (kp >= Lk || kp < g.lo_all) {
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
        if ((tid & 127) == 0) PF_TRACE(gi, 4 * i + 3);
        ++cnt; ++ci;
      }
      // ---- epilogue: read O back (and release it to the MMA warp), normalise, bf16, one 256-byte
      // row per thread
      uint32_t o[128];
      if (ci > 0) {
        mbar_wait(bar_o + 8 * i, oc & 1);
        tc_fence_after();
        tmem_ld32_nowait(tO + 0, *reinterpret_cast<uint32_t(*)[32]>(&o[0]));
        tmem_ld32_nowait(tO + 32, *reinterpret_cast<uint32_t(*)[32]>(&o[32]));
        tmem_ld32_nowait(tO + 64, *reinterpret_cast<uint32_t(*)[32]>(&o[64]));
        tmem_ld32_nowait(tO + 96, *reinterpret_cast<uint32_t(*)[32]>(&o[96]));
        tmem_wait_ld();
        tc_fence_before();
        mbar_arrive(bar_of + 8 * i);
        ++oc;
      } else {
#pragma unroll
        for (int j = 0; j < 128; ++j) o[j] = 0u;
      }
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
      float ssq = 0.f;
#if PF_OPT_EPI
      {
        // Per warp: 32 rows x 64 columns (4 KB, 16-byte chunks XOR-swizzled by row) staged in smem,
        // read back so that 8 lanes cover one 128-byte line: 4 full lines per store instruction
        // instead of 32 scattered 16-byte pieces (which kept the LSU -- and with it every mbarrier
        // poll of the other roles -- busy for thousands of cycles per item).
        uint8_t* stg = smem + PF_SO + warp * 4096;
        const int row_base = (warp & 3) * 32;
        __nv_bfloat16* dst0 = p.out + (g.o_row + i * PF_BM + row_base) * p.o_st + (int64_t)g.h * p.o_sh;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            const int j = 8 * half + c;
            uint4 v;
            v.x = pack_bf16(__uint_as_float(o[8 * j + 0]) * inv, __uint_as_float(o[8 * j + 1]) * inv);
            v.y = pack_bf16(__uint_as_float(o[8 * j + 2]) * inv, __uint_as_float(o[8 * j + 3]) * inv);
            v.z = pack_bf16(__uint_as_float(o[8 * j + 4]) * inv, __uint_as_float(o[8 * j + 5]) * inv);
            v.w = pack_bf16(__uint_as_float(o[8 * j + 6]) * inv, __uint_as_float(o[8 * j + 7]) * inv);
            if (p.sumsq) {   // the norm consumes the bf16 values
              const uint32_t* w4 = &v.x;
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                ssq = fmaf(bf16_lo(w4[e]), bf16_lo(w4[e]), ssq);
                ssq = fmaf(bf16_hi(w4[e]), bf16_hi(w4[e]), ssq);
              }
            }
            *reinterpret_cast<uint4*>(stg + lane * 128 + ((c ^ (lane & 7)) << 4)) = v;
          }
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int row = 4 * j + (lane >> 3), c = lane & 7;
            const uint4 v = *reinterpret_cast<const uint4*>(stg + row * 128 + ((c ^ (row & 7)) << 4));
            if (row_base + row < T.rows) stg_v4(dst0 + row * p.o_st + 64 * half + 8 * c, v);
          }
          __syncwarp();
        }
      }
#else
      if (r < T.rows) {
        __nv_bfloat16* dst = p.out + (g.o_row + i * PF_BM + r) * p.o_st + (int64_t)g.h * p.o_sh;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          uint4 v;
          v.x = pack_bf16(__uint_as_float(o[8 * j + 0]) * inv, __uint_as_float(o[8 * j + 1]) * inv);
          v.y = pack_bf16(__uint_as_float(o[8 * j + 2]) * inv, __uint_as_float(o[8 * j + 3]) * inv);
          v.z = pack_bf16(__uint_as_float(o[8 * j + 4]) * inv, __uint_as_float(o[8 * j + 5]) * inv);
          v.w = pack_bf16(__uint_as_float(o[8 * j + 6]) * inv, __uint_as_float(o[8 * j + 7]) * inv);
          stg_v4(dst + 8 * j, v);
        }
      }
#endif
      if (p.sumsq && r < T.rows) p.sumsq[(g.o_row + i * PF_BM + r) * p.H + g.h] = ssq;
      gt += g.n_steps;
    }
  } else {
    setmaxnreg_dec<PF_REG_LO>();   // warps 10-11 only complete the third warpgroup
  }
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

extern "C" int hymeta_swa_meta_prefill_varlen_sumsq(
    const void* q1, const void* k1, const void* v1, const void* q2, const void* meta_k,
    const void* meta_v, void* out, const int32_t* cu_q, const int32_t* cu_k,
    const int32_t* block_tables, int64_t bt_stride, int block_size, int N, int Tq, int Tk, int H,
    int Hkv, int D, int M, int max_seqlen_q, int window_left, float scale, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t q2_st, int64_t q2_sh,
    int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, int64_t o_st, int64_t o_sh,
    int64_t kc_sblk, int64_t vc_sblk, float* sumsq_out, void* stream) {
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
  p.sumsq = sumsq_out;

  static DynSmemAttr attr;
  attr.ensure(swa_prefill_kernel, PF_SMEM);
  p.n_z = p.N + (q2 ? 1 : 0);
  const int64_t n_items = (int64_t)H * p.n_qpairs * p.n_z;
  if (n_items > 0x7fffffff) return HYMETA_ERR_UNSUPPORTED;
  p.n_items = (int)n_items;
  dim3 grid((unsigned)std::min<int64_t>(n_items, sm_count()));
  swa_prefill_kernel<<<grid, PF_THREADS, PF_SMEM, (cudaStream_t)stream>>>(p, tmQ, tmK, tmV, tmQ2, tmMK, tmMV);
  count_launch();
  return check_launch();
}

extern "C" int hymeta_swa_meta_prefill_varlen(
    const void* q1, const void* k1, const void* v1, const void* q2, const void* meta_k,
    const void* meta_v, void* out, const int32_t* cu_q, const int32_t* cu_k,
    const int32_t* block_tables, int64_t bt_stride, int block_size, int N, int Tq, int Tk, int H,
    int Hkv, int D, int M, int max_seqlen_q, int window_left, float scale, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t q2_st, int64_t q2_sh,
    int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, int64_t o_st, int64_t o_sh,
    int64_t kc_sblk, int64_t vc_sblk, void* stream) {
  return hymeta_swa_meta_prefill_varlen_sumsq(q1, k1, v1, q2, meta_k, meta_v, out, cu_q, cu_k, block_tables,
                                              bt_stride, block_size, N, Tq, Tk, H, Hkv, D, M, max_seqlen_q,
                                              window_left, scale, q_st, q_sh, k_st, k_sh, v_st, v_sh, q2_st,
                                              q2_sh, mk_st, mk_sh, mv_st, mv_sh, o_st, o_sh, kc_sblk, vc_sblk,
                                              nullptr, stream);
}
