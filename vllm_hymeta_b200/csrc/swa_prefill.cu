// SWA + meta-prefix varlen prefill on tcgen05 / TMEM / TMA (a6/a7).
//
// One launch replaces the reference's two flash-attn launches (windowed causal pass + non-causal
// meta pass), the concat of q, the 6-kernel TorchScript LSE merge and the output concat
// (fa_kernel_metatoken.py:198-285): the meta keys are simply the first key tiles of every query
// tile's single online softmax.
//
// CTA = 128 query rows of one q-head (thread r <-> TMEM lane r <-> query row r), 128 threads,
// two CTAs per SM so one CTA's softmax overlaps the other's MMAs.
//   S[128 x 64]   = Q[128 x 128] K^T   : tcgen05.mma SS, K-major A/B (TMA, SWIZZLE_128B), fp32 in TMEM
//   P             = exp2(S*scale - m)  : registers; bf16 P written to smem (K-major canonical layout)
//   Ot[128 x 128] = P[128 x 64] V      : tcgen05.mma SS, V consumed MN-major exactly as TMA lands it
//   O             = O*alpha + Ot       : registers (one row per thread, no shuffles anywhere)
// K/V tiles (64 keys) are double-buffered; S(j+1) is issued behind PV(j) so the tensor pipe works
// while the CTA folds Ot(j).  grid.x = q-head so the 7 heads of a GQA group run back to back and
// share K/V tiles through L2.
#include "umma.cuh"

namespace hymeta {

constexpr int PF_BM = 128;   // query rows per CTA
constexpr int PF_BN = 64;    // keys per tile
constexpr int PF_SQ = 0;                       // [2][128][64] bf16
constexpr int PF_SK = 32768;                   // [2 stages][2][64][64]
constexpr int PF_SV = PF_SK + 32768;           // [2 stages][2][64][64]
constexpr int PF_SP = PF_SV + 32768;           // [128][64]
constexpr int PF_BAR = PF_SP + 16384;          // barriers + tmem ptr
constexpr int PF_SMEM = PF_BAR + 128;

struct SwaPrefillParams {
  const int32_t *cu_q, *cu_k;
  __nv_bfloat16* out;
  int N, H, Hkv, M, window_left, has_q2, n_qtiles;
  float scale_log2;
  int64_t o_st, o_sh;
};

__global__ void __launch_bounds__(128, 2)
swa_prefill_kernel(const SwaPrefillParams p, const __grid_constant__ CUtensorMap tmQ,
                   const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                   const __grid_constant__ CUtensorMap tmQ2, const __grid_constant__ CUtensorMap tmMK,
                   const __grid_constant__ CUtensorMap tmMV) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5;
  const int h = blockIdx.x, hkv = h / (p.H / p.Hkv);
  const int qt = p.n_qtiles - 1 - blockIdx.y;   // long tiles first
  const int n = blockIdx.z;
  const bool meta_q = n == p.N;                 // pseudo-sequence of the M meta queries

  int q_row0, Lq, Lk, k_row0;
  if (meta_q) { q_row0 = 0; Lq = p.M; Lk = 0; k_row0 = 0; }
  else {
    q_row0 = p.cu_q[n]; Lq = p.cu_q[n + 1] - q_row0;
    k_row0 = p.cu_k[n]; Lk = p.cu_k[n + 1] - k_row0;
  }
  if (qt * PF_BM >= Lq) return;

  const uint32_t sbase = smem_u32(smem);
  const uint32_t sQ = sbase + PF_SQ, sK = sbase + PF_SK, sV = sbase + PF_SV, sP = sbase + PF_SP;
  const uint32_t bar_q = sbase + PF_BAR, bar_kv0 = bar_q + 8, bar_s = bar_q + 24, bar_o = bar_q + 32;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + PF_BAR + 40);

  // key-position geometry of this query tile (queries aligned to the END of the keys)
  const int shift = Lk - Lq;
  const int rows = min(PF_BM, Lq - qt * PF_BM);
  const int pq_min = shift + qt * PF_BM, pq_max = pq_min + rows - 1;
  const int nm = (p.M + PF_BN - 1) / PF_BN;                       // meta tiles come first
  int kt_lo = 0, n_real = 0;
  if (!meta_q) {
    kt_lo = p.window_left < 0 ? 0 : max(0, pq_min - p.window_left) / PF_BN;
    n_real = pq_max >= 0 ? pq_max / PF_BN - kt_lo + 1 : 0;
  }
  const int nt = nm + n_real;

  if (tid == 0) {
    if (sbase & 1023) asm volatile("trap;");
    mbar_init(bar_q, 1); mbar_init(bar_kv0, 1); mbar_init(bar_kv0 + 8, 1);
    mbar_init(bar_s, 1); mbar_init(bar_o, 1);
    fence_mbar_init();
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
  }
  if (warp == 0) tmem_alloc<256>(smem_u32(tmem_slot));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tO = tmem + PF_BN;
  const uint32_t lane_off = (uint32_t)(warp * 32) << 16;

  auto load_kv = [&](int j) {   // thread 0 only
    const int st = j & 1;
    const uint32_t bar = bar_kv0 + 8 * st, dk = sK + st * 16384, dv = sV + st * 16384;
    mbar_expect_tx(bar, 32768);
    if (j < nm) {
      tma_load_3d(dk, &tmMK, bar, 0, hkv, j * PF_BN);  tma_load_3d(dk + 8192, &tmMK, bar, 64, hkv, j * PF_BN);
      tma_load_3d(dv, &tmMV, bar, 0, hkv, j * PF_BN);  tma_load_3d(dv + 8192, &tmMV, bar, 64, hkv, j * PF_BN);
    } else {
      const int row = k_row0 + (kt_lo + j - nm) * PF_BN;
      tma_load_3d(dk, &tmK, bar, 0, hkv, row);  tma_load_3d(dk + 8192, &tmK, bar, 64, hkv, row);
      tma_load_3d(dv, &tmV, bar, 0, hkv, row);  tma_load_3d(dv + 8192, &tmV, bar, 64, hkv, row);
    }
  };
  const uint32_t idesc_s = umma_idesc_bf16(PF_BM, PF_BN, 0, 0);
  const uint32_t idesc_o = umma_idesc_bf16(PF_BM, kD, 0, 1);
  auto issue_s = [&](int j) {   // thread 0 only: S = Q K^T
    const uint32_t kb = sK + (j & 1) * 16384;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const uint64_t da = umma_desc_sw128(sQ + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
      const uint64_t db = umma_desc_sw128(kb + (ks >> 2) * 8192 + (ks & 3) * 32, 16, 1024);
      umma_bf16(tS, da, db, idesc_s, ks > 0);
    }
    umma_commit(bar_s);
  };
  auto issue_pv = [&](int j) {  // thread 0 only: Ot = P V
    const uint32_t vb = sV + (j & 1) * 16384;
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      const uint64_t da = umma_desc_sw128(sP + ks * 32, 16, 1024);
      const uint64_t db = umma_desc_sw128(vb + ks * 2048, 8192, 1024);
      umma_bf16(tO, da, db, idesc_o, ks > 0);
    }
    umma_commit(bar_o);
  };

  if (tid == 0) {
    mbar_expect_tx(bar_q, 32768);
    const CUtensorMap* tq = meta_q ? &tmQ2 : &tmQ;
    tma_load_3d(sQ, tq, bar_q, 0, h, q_row0 + qt * PF_BM);
    tma_load_3d(sQ + 16384, tq, bar_q, 64, h, q_row0 + qt * PF_BM);
    load_kv(0);
    if (nt > 1) load_kv(1);
    mbar_wait(bar_q, 0);
    mbar_wait(bar_kv0, 0);
    tc_fence_after();
    issue_s(0);
  }

  const int r = tid;                    // my query row
  const int pq = pq_min + r;            // its key position
  float o_acc[kD];
#pragma unroll
  for (int i = 0; i < kD; ++i) o_acc[i] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;

  for (int j = 0; j < nt; ++j) {
    mbar_wait(bar_s, j & 1);
    tc_fence_after();
    float s[PF_BN];
    {
      float t0[32], t1[32];
      tmem_ld32(tS + lane_off, t0);
      tmem_ld32(tS + lane_off + 32, t1);
#pragma unroll
      for (int i = 0; i < 32; ++i) { s[i] = t0[i]; s[32 + i] = t1[i]; }
    }
    // ---- mask (tile-uniform decision) + scale
    bool need_mask;
    int kp0;
    if (j < nm) { kp0 = j * PF_BN; need_mask = kp0 + PF_BN > p.M; }
    else {
      kp0 = (kt_lo + j - nm) * PF_BN;
      need_mask = (kp0 + PF_BN - 1 > pq_min) || (kp0 + PF_BN > Lk) ||
                  (p.window_left >= 0 && kp0 < pq_max - p.window_left);
    }
    float mx = -INFINITY;
    if (need_mask) {
      int lo, hi;   // visible key positions for this row: [lo, hi]
      if (j < nm) { lo = 0; hi = p.M - 1; }
      else { hi = min(pq, Lk - 1); lo = p.window_left < 0 ? 0 : pq - p.window_left; }
#pragma unroll
      for (int i = 0; i < PF_BN; ++i) {
        const int kp = kp0 + i;
        s[i] = (kp >= lo && kp <= hi) ? s[i] * p.scale_log2 : -INFINITY;
        mx = fmaxf(mx, s[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < PF_BN; ++i) { s[i] *= p.scale_log2; mx = fmaxf(mx, s[i]); }
    }
    const float m_new = fmaxf(m_run, mx);
    const float alpha = (m_new == -INFINITY) ? 1.f : exp2f(m_run - m_new);
    const float msafe = (m_new == -INFINITY) ? 0.f : m_new;
    m_run = m_new;
    float rs = 0.f;
#pragma unroll
    for (int i = 0; i < PF_BN; ++i) { s[i] = exp2f(s[i] - msafe); rs += s[i]; }
    l_run = l_run * alpha + rs;
    // ---- P (bf16) -> smem, K-major SWIZZLE_128B row r
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      uint4 v;
      v.x = pack_bf16(s[8 * c + 0], s[8 * c + 1]); v.y = pack_bf16(s[8 * c + 2], s[8 * c + 3]);
      v.z = pack_bf16(s[8 * c + 4], s[8 * c + 5]); v.w = pack_bf16(s[8 * c + 6], s[8 * c + 7]);
      *reinterpret_cast<uint4*>(smem + PF_SP + sw128_chunk(r, c)) = v;
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      issue_pv(j);
      if (j + 1 < nt) {
        mbar_wait(bar_kv0 + 8 * ((j + 1) & 1), ((j + 1) >> 1) & 1);
        tc_fence_after();
        issue_s(j + 1);
      }
    }
    mbar_wait(bar_o, j & 1);
    tc_fence_after();
    if (tid == 0 && j + 2 < nt) load_kv(j + 2);   // stage j&1 is free: S(j) and PV(j) are complete
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float t[32];
      tmem_ld32(tO + lane_off + 32 * c, t);
#pragma unroll
      for (int i = 0; i < 32; ++i) o_acc[32 * c + i] = fmaf(o_acc[32 * c + i], alpha, t[i]);
    }
    tc_fence_before();   // orders these TMEM reads before the next iteration's barrier / MMAs
  }

  // ---- epilogue: normalise, bf16, one 256-byte row per thread
  if (r < rows) {
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    const int64_t orow = (meta_q ? 0 : (int64_t)(p.has_q2 ? p.M : 0) + q_row0) + qt * PF_BM + r;
    __nv_bfloat16* dst = p.out + orow * p.o_st + (int64_t)h * p.o_sh;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      uint4 v;
      v.x = pack_bf16(o_acc[8 * c + 0] * inv, o_acc[8 * c + 1] * inv);
      v.y = pack_bf16(o_acc[8 * c + 2] * inv, o_acc[8 * c + 3] * inv);
      v.z = pack_bf16(o_acc[8 * c + 4] * inv, o_acc[8 * c + 5] * inv);
      v.w = pack_bf16(o_acc[8 * c + 6] * inv, o_acc[8 * c + 7] * inv);
      stg_v4(dst + 8 * c, v);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(tmem);
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_swa_meta_prefill_varlen(
    const void* q1, const void* k1, const void* v1, const void* q2, const void* meta_k,
    const void* meta_v, void* out, const int32_t* cu_q, const int32_t* cu_k,
    const int32_t* block_tables, int64_t bt_stride, int block_size, int N, int Tq, int Tk, int H,
    int Hkv, int D, int M, int max_seqlen_q, int window_left, float scale, int64_t q_st, int64_t q_sh,
    int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t q2_st, int64_t q2_sh,
    int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, int64_t o_st, int64_t o_sh,
    int64_t kc_sblk, int64_t vc_sblk, void* stream) {
  (void)bt_stride; (void)block_size; (void)kc_sblk; (void)vc_sblk;
  if (N < 0 || Tq < 0 || Tk < 0 || H <= 0 || Hkv <= 0 || M < 0 || !out) return HYMETA_ERR_INVALID_ARG;
  if (D != kD || H % Hkv) return HYMETA_ERR_UNSUPPORTED;
  if (block_tables) return HYMETA_ERR_UNSUPPORTED;   // paged-prefix prefill: see DESIGN.md (next)
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
  int rc = HYMETA_OK;
  if (real) {
    if ((rc = mk3(&tmQ, q1, H, Tq, q_st, q_sh, PF_BM))) return rc;
    if ((rc = mk3(&tmK, k1, Hkv, Tk, k_st, k_sh, PF_BN))) return rc;
    if ((rc = mk3(&tmV, v1, Hkv, Tk, v_st, v_sh, PF_BN))) return rc;
  }
  if (M > 0) {
    if ((rc = mk3(&tmMK, meta_k, Hkv, M, mk_st, mk_sh, PF_BN))) return rc;
    if ((rc = mk3(&tmMV, meta_v, Hkv, M, mv_st, mv_sh, PF_BN))) return rc;
    if (q2 && (rc = mk3(&tmQ2, q2, H, M, q2_st, q2_sh, PF_BM))) return rc;
  }
  if (!real) { tmQ = tmQ2; tmK = tmMK; tmV = tmMV; }   // never dereferenced, but must be valid maps
  if (!q2) tmQ2 = tmQ;
  if (M == 0) { tmMK = tmK; tmMV = tmV; }

  SwaPrefillParams p;
  p.cu_q = cu_q; p.cu_k = cu_k; p.out = (__nv_bfloat16*)out;
  p.N = real ? N : 0; p.H = H; p.Hkv = Hkv; p.M = M; p.window_left = window_left;
  p.has_q2 = q2 ? 1 : 0;
  int msq = real ? max_seqlen_q : 0;
  if (q2 && M > msq) msq = M;
  if (msq <= 0) msq = real ? Tq : M;
  p.n_qtiles = (msq + PF_BM - 1) / PF_BM;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.o_st = o_st; p.o_sh = o_sh;

  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(swa_prefill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PF_SMEM);
    attr_set = true;
  }
  dim3 grid(H, p.n_qtiles, p.N + (q2 ? 1 : 0));
  swa_prefill_kernel<<<grid, 128, PF_SMEM, (cudaStream_t)stream>>>(p, tmQ, tmK, tmV, tmQ2, tmMK, tmMV);
  count_launch();
  return check_launch();
}
