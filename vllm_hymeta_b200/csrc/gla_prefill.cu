// GLA chunked prefill on tcgen05 / TMEM (a1/a3), varlen over cu_seqlens with slot
// states -- replaces the reference's per-request Python loop around fused_chunk_gla, its two host
// syncs per request and the 7x repeat_kv materialisation (modeling_hymeta.py:441-549).
//
// One CTA walks one (sequence, state-head) chain in chunks of C = 16 tokens.  The G = H/Hs q-heads
// sharing the state are stacked along the MMA M dimension: row = head*16 + token (G <= 8 -> M = 128).
// With b_t = cumsum(g) inside the chunk and the per-channel reference r = b_7 (mid-chunk, so every
// exponent is bounded by 8 steps of decay in either direction):
//     q~_t = q_t * scale * exp(b_t - r)      k~_s = k_s * exp(r - b_s)
//     A    = q~ k~^T  (causal, s <= t)               MMA  M128 N16  K128   (K-major x K-major)
//     O    = q~ (exp(r) (.) S)  +  A v               MMA  M128 N128 K128 + M128 N128 K16
//     dS   = k~^T v                                  MMA  M128 N128 K16    (MN-major x MN-major)
//     S   <- exp(b_C) (.) S + exp(b_C - r) (.) dS    fp32 master in registers (thread = d_k row)
// k~ is written to smem once and consumed both as the K-major B operand of A and as the MN-major A
// operand of dS (same bytes, two descriptors).  All accumulators live in TMEM (256 columns, two
// CTAs per SM); the bf16 copy of the decayed state is rewritten to smem each chunk as the MN-major
// B operand of O.
#include "umma.cuh"

namespace hymeta {

constexpr int GC = 16;                          // tokens per chunk
constexpr int GP_SQ = 0;                        // q~  [2][128][64] bf16            32 KB
constexpr int GP_SS = 32768;                    // S   [2][128 dk][64 dv] bf16      32 KB
constexpr int GP_SP = 65536;                    // P   [128][64] bf16 (K=16 used)   16 KB
constexpr int GP_SK = 81920;                    // k~  [2][16][64] bf16              4 KB
constexpr int GP_SV = 86016;                    // v   [2][16][64] bf16              4 KB
constexpr int GP_EQ = 90112;                    // scale*exp(b_t - r) [16][128] f32  8 KB
constexpr int GP_BAR = 98304;
constexpr int GP_SMEM = GP_BAR + 64;

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

template <bool GF32>
__global__ void __launch_bounds__(128, 2) gla_prefill_kernel(const GlaPrefillParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5;
  const int hs = blockIdx.x, n = blockIdx.y;
  const int G = p.H / p.Hs, h0 = hs * G, hk = hs / (p.Hs / p.Hk), hv = hs / (p.Hs / p.Hv);
  const int tok_begin = p.cu[n], len = p.cu[n + 1] - tok_begin;
  const int slot = p.slot[n], mode = p.init_mode[n];

  const uint32_t sbase = smem_u32(smem);
  const uint32_t sQ = sbase + GP_SQ, sS = sbase + GP_SS, sP = sbase + GP_SP, sK = sbase + GP_SK,
                 sV = sbase + GP_SV;
  float* sEq = reinterpret_cast<float*>(smem + GP_EQ);
  const uint32_t bar1 = sbase + GP_BAR, bar2 = bar1 + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + GP_BAR + 16);

  // ---- fp32 master state: thread = d_k row, 128 d_v columns in registers
  float S[kD];
  {
    const __nv_bfloat16* src = nullptr;
    if (mode == 1 && slot >= 0) src = p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh + tid * kD;
    else if (mode == 2 && p.meta_state) src = p.meta_state + (int64_t)hs * p.meta_sh + tid * kD;
    if (src) {
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const uint4 t = ldg_v4(src + 8 * c);
        S[8 * c + 0] = bf16_lo(t.x); S[8 * c + 1] = bf16_hi(t.x); S[8 * c + 2] = bf16_lo(t.y);
        S[8 * c + 3] = bf16_hi(t.y); S[8 * c + 4] = bf16_lo(t.z); S[8 * c + 5] = bf16_hi(t.z);
        S[8 * c + 6] = bf16_lo(t.w); S[8 * c + 7] = bf16_hi(t.w);
      }
    } else {
#pragma unroll
      for (int i = 0; i < kD; ++i) S[i] = 0.f;
    }
  }

  if (tid == 0) {
    if (sbase & 1023) asm volatile("trap;");
    mbar_init(bar1, 1); mbar_init(bar2, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<256>(smem_u32(tmem_slot));
  // zero P once: only its first 32 bytes per row are ever rewritten / read by the K=16 MMA
  for (int i = tid; i < 16384 / 16; i += 128) *reinterpret_cast<uint4*>(smem + GP_SP + i * 16) = make_uint4(0, 0, 0, 0);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tO = tmem, tS = tmem + 128, tA = tmem + 128;
  const uint32_t lane_off = (uint32_t)(warp * 32) << 16;

  const uint32_t idesc_a = umma_idesc_bf16(128, 16, 0, 0);
  const uint32_t idesc_o = umma_idesc_bf16(128, 128, 0, 1);
  const uint32_t idesc_s = umma_idesc_bf16(128, 128, 1, 1);

  // row-owner view of this thread (O / A rows): head slot hh, token t
  const int r_hh = tid >> 4, r_t = tid & 15;
  // q~ producer view: fixed 8-channel chunk, two tokens, all head slots
  const int qc = tid & 15, qt0 = tid >> 4;

  const int n_chunks = (len + GC - 1) / GC;
  for (int ch = 0; ch < n_chunks; ++ch) {
    const int tok0 = tok_begin + ch * GC;
    const int clen = min(GC, len - ch * GC);

    // (a) gates for channel d_k = tid: b_t, r = b_7, exponentials
    float bcum[GC];
    {
      float acc = 0.f;
#pragma unroll
      for (int t = 0; t < GC; ++t) {
        float gv = 0.f;
        if (t < clen) {
          const int64_t off = (int64_t)(tok0 + t) * p.g_st + (int64_t)hk * p.g_sh + tid;
          gv = GF32 ? reinterpret_cast<const float*>(p.g)[off]
                    : __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.g)[off]);
        }
        acc += gv;
        bcum[t] = acc;
      }
    }
    const float rref = bcum[GC / 2 - 1];
    const float e_r = __expf(rref), e_c = __expf(bcum[GC - 1]), e_cr = __expf(bcum[GC - 1] - rref);
    // (b) k~ for this channel, all tokens -> smem (K-major B of A  ==  MN-major A of dS)
#pragma unroll
    for (int t = 0; t < GC; ++t) {
      const float d = bcum[t] - rref;
      sEq[t * kD + tid] = p.scale * __expf(d);
      float kv = 0.f;
      if (t < clen) kv = __bfloat162float(p.k[(int64_t)(tok0 + t) * p.k_st + (int64_t)hk * p.k_sh + tid]);
      *reinterpret_cast<__nv_bfloat16*>(smem + GP_SK + (tid >> 6) * 2048 + sw128_off(t, tid & 63)) =
          __float2bfloat16_rn(kv * __expf(-d));
    }
    // (e) bf16 copy of exp(r) (.) S, row d_k = tid (MN-major B of O)
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      uint4 v4;
      v4.x = pack_bf16(S[8 * c + 0] * e_r, S[8 * c + 1] * e_r); v4.y = pack_bf16(S[8 * c + 2] * e_r, S[8 * c + 3] * e_r);
      v4.z = pack_bf16(S[8 * c + 4] * e_r, S[8 * c + 5] * e_r); v4.w = pack_bf16(S[8 * c + 6] * e_r, S[8 * c + 7] * e_r);
      *reinterpret_cast<uint4*>(smem + GP_SS + (c >> 3) * 16384 + sw128_chunk(tid, c & 7)) = v4;
    }
    // (c) v tile: 16 tokens x 16 chunks of 8 channels = 256 items
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * 128, t = idx >> 4, c = idx & 15;
      uint4 v4 = make_uint4(0, 0, 0, 0);
      if (t < clen) v4 = ldg_nc_v4(p.v + (int64_t)(tok0 + t) * p.v_st + (int64_t)hv * p.v_sh + c * 8);
      *reinterpret_cast<uint4*>(smem + GP_SV + (c >> 3) * 2048 + sw128_chunk(t, c & 7)) = v4;
    }
    __syncthreads();   // sEq visible
    // (d) q~ tile: thread owns channels [8qc, 8qc+8) of tokens qt0 and qt0+8 for every head slot
    {
      float e0[8], e1[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { e0[j] = sEq[qt0 * kD + qc * 8 + j]; e1[j] = sEq[(qt0 + 8) * kD + qc * 8 + j]; }
#pragma unroll
      for (int it = 0; it < 16; ++it) {
        const int hh = it >> 1, t = qt0 + (it & 1) * 8, row = hh * 16 + t;
        uint4 v4 = make_uint4(0, 0, 0, 0);
        if (hh < G && t < clen)
          v4 = ldg_nc_v4(p.q + (int64_t)(tok0 + t) * p.q_st + (int64_t)(h0 + hh) * p.q_sh + qc * 8);
        const float* e = (it & 1) ? e1 : e0;
        uint4 w;
        w.x = pack_bf16(bf16_lo(v4.x) * e[0], bf16_hi(v4.x) * e[1]);
        w.y = pack_bf16(bf16_lo(v4.y) * e[2], bf16_hi(v4.y) * e[3]);
        w.z = pack_bf16(bf16_lo(v4.z) * e[4], bf16_hi(v4.z) * e[5]);
        w.w = pack_bf16(bf16_lo(v4.w) * e[6], bf16_hi(v4.w) * e[7]);
        *reinterpret_cast<uint4*>(smem + GP_SQ + (qc >> 3) * 16384 + sw128_chunk(row, qc & 7)) = w;
      }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {   // A = q~ k~^T
        const uint64_t da = umma_desc_sw128(sQ + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
        const uint64_t db = umma_desc_sw128(sK + (ks >> 2) * 2048 + (ks & 3) * 32, 16, 1024);
        umma_bf16(tA, da, db, idesc_a, ks > 0);
      }
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) {   // O = q~ (exp(r) S)
        const uint64_t da = umma_desc_sw128(sQ + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
        const uint64_t db = umma_desc_sw128(sS + ks * 2048, 16384, 1024);
        umma_bf16(tO, da, db, idesc_o, ks > 0);
      }
      umma_commit(bar1);
    }
    mbar_wait(bar1, ch & 1);
    tc_fence_after();
    {   // causal A -> bf16 P (row owner), first 32 bytes of the SWIZZLE_128B row
      float a[16];
      tmem_ld16(tA + lane_off, a);
#pragma unroll
      for (int s = 0; s < 16; ++s) a[s] = (s <= r_t) ? a[s] : 0.f;
      uint4 w0, w1;
      w0.x = pack_bf16(a[0], a[1]); w0.y = pack_bf16(a[2], a[3]); w0.z = pack_bf16(a[4], a[5]); w0.w = pack_bf16(a[6], a[7]);
      w1.x = pack_bf16(a[8], a[9]); w1.y = pack_bf16(a[10], a[11]); w1.z = pack_bf16(a[12], a[13]); w1.w = pack_bf16(a[14], a[15]);
      *reinterpret_cast<uint4*>(smem + GP_SP + sw128_chunk(tid, 0)) = w0;
      *reinterpret_cast<uint4*>(smem + GP_SP + sw128_chunk(tid, 1)) = w1;
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      umma_bf16(tO, umma_desc_sw128(sP, 16, 1024), umma_desc_sw128(sV, 2048, 1024), idesc_o, 1);     // O += P v
      umma_bf16(tS, umma_desc_sw128(sK, 2048, 1024), umma_desc_sw128(sV, 2048, 1024), idesc_s, 0);   // dS = k~^T v
      umma_commit(bar2);
    }
    mbar_wait(bar2, ch & 1);
    tc_fence_after();
    // state fold (d_k row owner)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float t[32];
      tmem_ld32(tS + lane_off + 32 * c, t);
#pragma unroll
      for (int i = 0; i < 32; ++i) S[32 * c + i] = S[32 * c + i] * e_c + t[i] * e_cr;
    }
    // output rows (row owner): 256 contiguous bytes per (token, head)
    {
      const bool live = r_hh < G && r_t < clen;
      __nv_bfloat16* dst = p.o + (int64_t)(tok0 + r_t) * p.o_st + (int64_t)(h0 + r_hh) * p.o_sh;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float t[32];
        tmem_ld32(tO + lane_off + 32 * c, t);
        if (live) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 w;
            w.x = pack_bf16(t[8 * j + 0], t[8 * j + 1]); w.y = pack_bf16(t[8 * j + 2], t[8 * j + 3]);
            w.z = pack_bf16(t[8 * j + 4], t[8 * j + 5]); w.w = pack_bf16(t[8 * j + 6], t[8 * j + 7]);
            stg_v4(dst + 32 * c + 8 * j, w);
          }
        }
      }
    }
    tc_fence_before();
  }

  if (p.write_state && slot >= 0) {
    __nv_bfloat16* dst = p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh + tid * kD;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      uint4 w;
      w.x = pack_bf16(S[8 * c + 0], S[8 * c + 1]); w.y = pack_bf16(S[8 * c + 2], S[8 * c + 3]);
      w.z = pack_bf16(S[8 * c + 4], S[8 * c + 5]); w.w = pack_bf16(S[8 * c + 6], S[8 * c + 7]);
      stg_v4(dst + 8 * c, w);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(tmem);
}

}  // namespace hymeta

using namespace hymeta;

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
  if (g_is_fp32) gla_prefill_kernel<true><<<grid, 128, GP_SMEM, (cudaStream_t)stream>>>(p);
  else gla_prefill_kernel<false><<<grid, 128, GP_SMEM, (cudaStream_t)stream>>>(p);
  count_launch();
  return check_launch();
}
