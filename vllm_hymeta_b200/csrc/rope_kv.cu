// Fused SWA prologue (f3): NeoX rotary embedding of q and k, the paged-KV write of the
// rotated k and of v, and the meta-K/V capture of the first pass -- one launch, one read of the
// qkv_proj output, instead of ops.rotary_embedding + reshape_and_cache_flash + two copy_ launches
// (modeling_hymeta.py:220-234,762,776-792,1314-1324; meta_attention.py:790-799).
//
// Arithmetic follows the reference's rotary op instantiated for bf16 (cache_dtype=torch.bfloat16,
// modeling_hymeta.py:1187): every product and the final add/sub are rounded to bf16
//   x' = bf16(bf16(x*cos) - bf16(y*sin)),  y' = bf16(bf16(y*cos) + bf16(x*sin)),
// x = head[0:D/2], y = head[D/2:D], cos/sin read from the caller's bf16 table [max_pos, D]
// (cos | sin).  The product of two bf16 values is exact in fp32, so one rounding of the fp32 product
// is the bf16 product.  HBM-bound; compiled without fast-math (no FTZ) so it is bit-exact.
#include "common.cuh"

namespace hymeta {

struct RopeKvParams {
  const __nv_bfloat16 *q_in, *k_in, *v_in, *cos_sin;
  __nv_bfloat16 *q_out, *k_out, *k_cache, *v_cache, *meta_k, *meta_v;
  const int64_t *positions, *slot_mapping;
  int T, H, Hkv, n_meta, block_size, max_pos, fp32_math;
  int64_t pos_offset;
  int64_t qi_st, qi_sh, ki_st, ki_sh, vi_st, vi_sh, qo_st, qo_sh, ko_st, ko_sh;
  int64_t kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh, mk_st, mk_sh, mv_st, mv_sh;
};

// FP32: the products (exact in fp32) are combined in fp32 and rounded to bf16 ONCE -- the arithmetic of
// the rotary op in the vLLM 0.22 installed next to this package (bit-identical to it on B200:
// tests/test_gpu_live_libs.py).  !FP32: rounding after every operation (vLLM 0.10.0 source).
template <bool FP32>
__device__ __forceinline__ void rope8(const uint4& x, const uint4& y, const uint4& c,
                                      const uint4& s, uint4& xo, uint4& yo) {
  const uint32_t* xp = &x.x; const uint32_t* yp = &y.x;
  const uint32_t* cp = &c.x; const uint32_t* sp = &s.x;
  uint32_t* xop = &xo.x; uint32_t* yop = &yo.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float x0 = bf16_lo(xp[i]), x1 = bf16_hi(xp[i]), y0 = bf16_lo(yp[i]), y1 = bf16_hi(yp[i]);
    const float c0 = bf16_lo(cp[i]), c1 = bf16_hi(cp[i]), s0 = bf16_lo(sp[i]), s1 = bf16_hi(sp[i]);
    if (FP32) {
      xop[i] = pack_bf16(__fsub_rn(__fmul_rn(x0, c0), __fmul_rn(y0, s0)),
                         __fsub_rn(__fmul_rn(x1, c1), __fmul_rn(y1, s1)));
      yop[i] = pack_bf16(__fadd_rn(__fmul_rn(y0, c0), __fmul_rn(x0, s0)),
                         __fadd_rn(__fmul_rn(y1, c1), __fmul_rn(x1, s1)));
      continue;
    }
    xop[i] = pack_bf16(__fsub_rn(round_bf16(__fmul_rn(x0, c0)), round_bf16(__fmul_rn(y0, s0))),
                       __fsub_rn(round_bf16(__fmul_rn(x1, c1)), round_bf16(__fmul_rn(y1, s1))));
    yop[i] = pack_bf16(__fadd_rn(round_bf16(__fmul_rn(y0, c0)), round_bf16(__fmul_rn(x0, s0))),
                       __fadd_rn(round_bf16(__fmul_rn(y1, c1)), round_bf16(__fmul_rn(x1, s1))));
  }
}

// One work item = (token, head of [q | k | v], 16-byte chunk c of the first half) and the matching
// chunk of the second half: 2 x 16 B in, 2 x 16 B out (x2 for k: k_out and the page).
// IdxT = uint32_t whenever the item count fits (always in practice: T < 7.4 M tokens at 36 heads):
// the per-item div/mod by the head count is then 32-bit instead of emulated 64-bit arithmetic.
template <typename IdxT>
__global__ void __launch_bounds__(256) rope_kv_write_kernel(const RopeKvParams p) {
  constexpr int HC = kD / 16;  // chunks per half row (8)
  pdl_wait();
  pdl_launch_dependents();
  const int heads = p.H + 2 * p.Hkv;
  const IdxT total = (IdxT)p.T * (IdxT)heads * HC;
  const IdxT stride = (IdxT)gridDim.x * blockDim.x;
  for (IdxT i = (IdxT)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int c = (int)(i % HC);
    const int hh = (int)((i / HC) % (IdxT)heads);
    const int64_t t = (int64_t)(i / ((IdxT)HC * (IdxT)heads));
    const bool is_meta = t < p.n_meta;
    if (hh >= p.H + p.Hkv) {  // value head: plain copy into the page / meta cache
      const int h = hh - p.H - p.Hkv;
      __nv_bfloat16* dst;
      if (is_meta) {
        dst = p.meta_v + t * p.mv_st + h * p.mv_sh;
      } else {
        if (!p.v_cache) continue;
        const int64_t slot = p.slot_mapping[t - p.n_meta];
        if (slot < 0) continue;
        dst = p.v_cache + (slot / p.block_size) * p.vc_sblk + (slot % p.block_size) * p.vc_stok +
              h * p.vc_sh;
      }
      const __nv_bfloat16* src = p.v_in + t * p.vi_st + h * p.vi_sh;
      const uint4 a = ldg_nc_v4(src + c * 8), b = ldg_nc_v4(src + kD / 2 + c * 8);
      stg_v4(dst + c * 8, a);
      stg_v4(dst + kD / 2 + c * 8, b);
      continue;
    }
    int64_t pos = is_meta ? (int64_t)t : p.positions[t - p.n_meta] + p.pos_offset;
    pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);
    const __nv_bfloat16* cs = p.cos_sin + pos * kD;
    const uint4 cv = ldg_nc_v4(cs + c * 8), sv = ldg_nc_v4(cs + kD / 2 + c * 8);
    const bool is_q = hh < p.H;
    const int h = is_q ? hh : hh - p.H;
    const __nv_bfloat16* src = is_q ? p.q_in + t * p.qi_st + h * p.qi_sh
                                    : p.k_in + t * p.ki_st + h * p.ki_sh;
    // coherent loads: q_out/k_out may alias q_in/k_in (in-place, as the reference op)
    const uint4 x = ldg_v4(src + c * 8), y = ldg_v4(src + kD / 2 + c * 8);
    uint4 xo, yo;
    if (p.fp32_math) rope8<true>(x, y, cv, sv, xo, yo);
    else rope8<false>(x, y, cv, sv, xo, yo);
    __nv_bfloat16* out = is_q ? p.q_out + t * p.qo_st + h * p.qo_sh
                              : p.k_out + t * p.ko_st + h * p.ko_sh;
    stg_v4(out + c * 8, xo);
    stg_v4(out + kD / 2 + c * 8, yo);
    if (is_q) continue;
    __nv_bfloat16* dst;
    if (is_meta) {
      dst = p.meta_k + t * p.mk_st + h * p.mk_sh;
    } else {
      if (!p.k_cache) continue;
      const int64_t slot = p.slot_mapping[t - p.n_meta];
      if (slot < 0) continue;
      dst = p.k_cache + (slot / p.block_size) * p.kc_sblk + (slot % p.block_size) * p.kc_stok +
            h * p.kc_sh;
    }
    stg_v4(dst + c * 8, xo);
    stg_v4(dst + kD / 2 + c * 8, yo);
  }
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_rope_kv_write_paged(
    const void* q_in, const void* k_in, const void* v_in, void* q_out, void* k_out,
    const void* cos_sin_cache, int max_pos, const int64_t* positions, int64_t pos_offset,
    void* k_cache, void* v_cache, const int64_t* slot_mapping, void* meta_k, void* meta_v,
    int n_meta, int T, int H, int Hkv, int D, int block_size,
    int64_t qi_st, int64_t qi_sh, int64_t ki_st, int64_t ki_sh, int64_t vi_st, int64_t vi_sh,
    int64_t qo_st, int64_t qo_sh, int64_t ko_st, int64_t ko_sh,
    int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk, int64_t vc_stok,
    int64_t vc_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, int fp32_math,
    void* stream) {
  if (T == 0) return HYMETA_OK;
  if (!q_in || !k_in || !v_in || !q_out || !k_out || !cos_sin_cache || T < 0 || H <= 0 ||
      Hkv <= 0 || n_meta < 0 || n_meta > T || max_pos <= 0)
    return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  if (n_meta > 0 && (!meta_k || !meta_v)) return HYMETA_ERR_INVALID_ARG;
  if (T > n_meta && !positions) return HYMETA_ERR_INVALID_ARG;
  const bool paged = k_cache != nullptr;  // null caches: profiling run (kv_cache.numel() == 0)
  if (paged && (!v_cache || (!slot_mapping && T > n_meta) || block_size <= 0))
    return HYMETA_ERR_INVALID_ARG;
  if (!paged) v_cache = nullptr;
  const int64_t strides[] = {qi_st, qi_sh, ki_st, ki_sh, vi_st, vi_sh, qo_st, qo_sh, ko_st, ko_sh,
                             kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh,
                             mk_st, mk_sh, mv_st, mv_sh};
  for (int64_t s : strides)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;  // 16-byte vector access
  RopeKvParams p;
  p.q_in = (const __nv_bfloat16*)q_in; p.k_in = (const __nv_bfloat16*)k_in;
  p.v_in = (const __nv_bfloat16*)v_in; p.cos_sin = (const __nv_bfloat16*)cos_sin_cache;
  p.q_out = (__nv_bfloat16*)q_out; p.k_out = (__nv_bfloat16*)k_out;
  p.k_cache = (__nv_bfloat16*)k_cache; p.v_cache = (__nv_bfloat16*)v_cache;
  p.meta_k = (__nv_bfloat16*)meta_k; p.meta_v = (__nv_bfloat16*)meta_v;
  p.positions = positions; p.slot_mapping = slot_mapping;
  p.T = T; p.H = H; p.Hkv = Hkv; p.n_meta = n_meta; p.block_size = paged ? block_size : 1;
  p.max_pos = max_pos; p.pos_offset = pos_offset; p.fp32_math = fp32_math ? 1 : 0;
  p.qi_st = qi_st; p.qi_sh = qi_sh; p.ki_st = ki_st; p.ki_sh = ki_sh; p.vi_st = vi_st;
  p.vi_sh = vi_sh; p.qo_st = qo_st; p.qo_sh = qo_sh; p.ko_st = ko_st; p.ko_sh = ko_sh;
  p.kc_sblk = kc_sblk; p.kc_stok = kc_stok; p.kc_sh = kc_sh; p.vc_sblk = vc_sblk;
  p.vc_stok = vc_stok; p.vc_sh = vc_sh; p.mk_st = mk_st; p.mk_sh = mk_sh; p.mv_st = mv_st;
  p.mv_sh = mv_sh;
  const int64_t total = (int64_t)T * (H + 2 * Hkv) * (kD / 16);
  const int64_t want = (total + 255) / 256;
  const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
  // the grid-stride loop adds `stride` past `total`: keep head-room below 2^32 for the 32-bit form
  if (total < (int64_t)0xf0000000LL)
    launch_pdl(rope_kv_write_kernel<uint32_t>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, p);
  else
    launch_pdl(rope_kv_write_kernel<unsigned long long>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, p);
  count_launch();
  return check_launch();
}
