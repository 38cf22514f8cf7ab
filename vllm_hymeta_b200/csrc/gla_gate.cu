// GLA feature map + gate pre-processing (a4; 8f: f2) -- one elementwise pass that replaces
// the ~10 eager torch launches of the reference between qkv_proj and the GLA kernels:
//   70B (modeling_hymeta.py:611-629):  q = relu(q);  s = sigmoid(float(k));
//        no lower bound / first layer of a group:  k' = min(s, clamp_max);  g = log(1 - k')
//        else:                                     g' = 1 - (1 - lb) * s;   k' = 1 - g';  g = log(g')
//        k' -> bf16, g stays fp32
//   7B  (gla_attention.py:251-262):    q = relu(q);  k = relu(k);  g = bf16(logsigmoid(gk)) / 16
// It reads q / k / gk as strided views of the projection output (no split / reshape copies) and
// never materialises repeat_kv: k and g are produced per INPUT k-head.  With a lower bound the
// reference's lb is indexed by q-head (:620-625), so k' and g then have H heads (each q-head of a
// GQA group gets its own gate) -- the kernel broadcasts the raw k-head across the group itself.
// This file is compiled WITHOUT --use_fast_math (see __graft_entry__.py): expf/logf/div are the
// IEEE-accurate ones so g matches the fp32 torch reference to the last few ulps.
#include "common.cuh"

namespace hymeta {

struct GateParams {
  const __nv_bfloat16 *q_in, *k_in, *gk_in;
  __nv_bfloat16 *q_out, *k_out;
  void* g_out;
  const float* lb;   // [Ho, D] or null
  int T, H, Hk, Ho;  // Ho = heads of k_out / g_out (Hk, or H when lb broadcasts over q-heads)
  int64_t qi_st, qi_sh, ki_st, ki_sh, gi_st, gi_sh, qo_st, qo_sh, ko_st, ko_sh, go_st, go_sh;
  float clamp_max, norm;
};

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 v;
  v.x = pack_bf16(f[0], f[1]); v.y = pack_bf16(f[2], f[3]);
  v.z = pack_bf16(f[4], f[5]); v.w = pack_bf16(f[6], f[7]);
  return v;
}
__device__ __forceinline__ uint4 relu8(const uint4& v) {
  // bf16 relu on the packed words: clear a half whose sign bit is set (-0 -> +0 like torch.relu;
  // NaN payloads with the sign bit set do not occur in projection outputs)
  auto r = [](uint32_t w) {
    const uint32_t neg = (w >> 15) & 0x00010001u;       // sign bits -> bit 0 of each half
    return w & ~(neg * 0xffffu);
  };
  return make_uint4(r(v.x), r(v.y), r(v.z), r(v.w));
}

// MODE 0: 70B without lower bound, 1: 70B with lower bound, 2: 7B, 3: 7B gate only (g from gk; relu of
// q and k is left to the GLA kernels, which apply it on the fly to the raw projection outputs)
template <int MODE>
__global__ void __launch_bounds__(256) gla_gate_kernel(const GateParams p) {
  constexpr int CPR = kD / 8;   // 16-byte chunks per head row
  const int64_t nq = (int64_t)p.T * p.H * CPR, nk = (int64_t)p.T * p.Ho * CPR;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nq + nk;
       i += (int64_t)gridDim.x * blockDim.x) {
    if (i < nq) {
      const int c = (int)(i % CPR), h = (int)((i / CPR) % p.H);
      const int64_t t = i / ((int64_t)CPR * p.H);
      const uint4 v = ldg_nc_v4(p.q_in + t * p.qi_st + h * p.qi_sh + c * 8);
      stg_v4(p.q_out + t * p.qo_st + h * p.qo_sh + c * 8, relu8(v));
      continue;
    }
    const int64_t j = i - nq;
    const int c = (int)(j % CPR), ho = (int)((j / CPR) % p.Ho);
    const int64_t t = j / ((int64_t)CPR * p.Ho);
    const int hk = ho / (p.Ho / p.Hk);
    uint4 kraw = make_uint4(0, 0, 0, 0);
    if (MODE != 3) kraw = ldg_nc_v4(p.k_in + t * p.ki_st + hk * p.ki_sh + c * 8);
    if (MODE >= 2) {
      if (MODE == 2) stg_v4(p.k_out + t * p.ko_st + ho * p.ko_sh + c * 8, relu8(kraw));
      float x[8], g[8];
      unpack8(ldg_nc_v4(p.gk_in + t * p.gi_st + hk * p.gi_sh + c * 8), x);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        // logsigmoid(x) = min(x, 0) - log1p(exp(-|x|))   (ATen's formulation), rounded to bf16
        // BEFORE the division, as the bf16 torch path does
        const float ls = fminf(x[e], 0.f) - log1pf(expf(-fabsf(x[e])));
        g[e] = round_bf16(ls) / p.norm;
      }
      stg_v4(reinterpret_cast<__nv_bfloat16*>(p.g_out) + t * p.go_st + ho * p.go_sh + c * 8, pack8(g));
    } else {
      float x[8], kk[8], g[8];
      unpack8(kraw, x);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float s = 1.f / (1.f + expf(-x[e]));
        if (MODE == 0) {
          kk[e] = fminf(s, p.clamp_max);
          g[e] = logf(1.f - kk[e]);
        } else {
          const float lb = p.lb[ho * kD + c * 8 + e];
          const float gp = 1.f - (1.f - lb) * s;
          kk[e] = 1.f - gp;
          g[e] = logf(gp);
        }
      }
      stg_v4(p.k_out + t * p.ko_st + ho * p.ko_sh + c * 8, pack8(kk));
      float* gd = reinterpret_cast<float*>(p.g_out) + t * p.go_st + ho * p.go_sh + c * 8;
      *reinterpret_cast<float4*>(gd) = make_float4(g[0], g[1], g[2], g[3]);
      *reinterpret_cast<float4*>(gd + 4) = make_float4(g[4], g[5], g[6], g[7]);
    }
  }
}

static int launch_gate(const GateParams& p, int mode, cudaStream_t st) {
  const int64_t total = ((int64_t)p.T * (p.H + p.Ho)) * (kD / 8);
  if (total == 0) return HYMETA_OK;
  const int64_t want = (total + 255) / 256;
  const int blocks = (int)(want < 148 * 16 ? want : 148 * 16);
  if (mode == 0) gla_gate_kernel<0><<<blocks, 256, 0, st>>>(p);
  else if (mode == 1) gla_gate_kernel<1><<<blocks, 256, 0, st>>>(p);
  else if (mode == 2) gla_gate_kernel<2><<<blocks, 256, 0, st>>>(p);
  else gla_gate_kernel<3><<<blocks, 256, 0, st>>>(p);
  count_launch();
  return check_launch();
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_gla_gate_70b(const void* q_in, const void* k_in, void* q_out, void* k_out,
                                   float* g_out, const float* lower_bound, int T, int H, int Hk,
                                   int Ho, int D, int64_t qi_st, int64_t qi_sh, int64_t ki_st,
                                   int64_t ki_sh, int64_t qo_st, int64_t qo_sh, int64_t ko_st,
                                   int64_t ko_sh, int64_t go_st, int64_t go_sh, float clamp_max,
                                   void* stream) {
  if (T == 0) return HYMETA_OK;
  if (!q_in || !k_in || !q_out || !k_out || !g_out || T < 0 || H <= 0 || Hk <= 0 || Ho <= 0)
    return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  if (Ho % Hk || (!lower_bound && Ho != Hk)) return HYMETA_ERR_INVALID_ARG;
  const int64_t s8[] = {qi_st, qi_sh, ki_st, ki_sh, qo_st, qo_sh, ko_st, ko_sh};
  for (int64_t s : s8)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;
  if ((go_st % 4) || (go_sh % 4)) return HYMETA_ERR_INVALID_ARG;
  GateParams p{};
  p.q_in = (const __nv_bfloat16*)q_in; p.k_in = (const __nv_bfloat16*)k_in; p.gk_in = nullptr;
  p.q_out = (__nv_bfloat16*)q_out; p.k_out = (__nv_bfloat16*)k_out; p.g_out = g_out;
  p.lb = lower_bound; p.T = T; p.H = H; p.Hk = Hk; p.Ho = Ho;
  p.qi_st = qi_st; p.qi_sh = qi_sh; p.ki_st = ki_st; p.ki_sh = ki_sh; p.qo_st = qo_st; p.qo_sh = qo_sh;
  p.ko_st = ko_st; p.ko_sh = ko_sh; p.go_st = go_st; p.go_sh = go_sh;
  p.clamp_max = clamp_max; p.norm = 1.f;
  return launch_gate(p, lower_bound ? 1 : 0, (cudaStream_t)stream);
}

extern "C" int hymeta_gla_gate_7b(const void* q_in, const void* k_in, const void* gk_in, void* q_out,
                                  void* k_out, void* g_out, int T, int H, int Hk, int D,
                                  int64_t qi_st, int64_t qi_sh, int64_t ki_st, int64_t ki_sh,
                                  int64_t gi_st, int64_t gi_sh, int64_t qo_st, int64_t qo_sh,
                                  int64_t ko_st, int64_t ko_sh, int64_t go_st, int64_t go_sh,
                                  float normalizer, void* stream) {
  if (T == 0) return HYMETA_OK;
  if (!q_in || !k_in || !gk_in || !q_out || !k_out || !g_out || T < 0 || H <= 0 || Hk <= 0 ||
      normalizer == 0.f)
    return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  const int64_t s8[] = {qi_st, qi_sh, ki_st, ki_sh, gi_st, gi_sh, qo_st, qo_sh, ko_st, ko_sh, go_st, go_sh};
  for (int64_t s : s8)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;
  GateParams p{};
  p.q_in = (const __nv_bfloat16*)q_in; p.k_in = (const __nv_bfloat16*)k_in;
  p.gk_in = (const __nv_bfloat16*)gk_in;
  p.q_out = (__nv_bfloat16*)q_out; p.k_out = (__nv_bfloat16*)k_out; p.g_out = g_out;
  p.lb = nullptr; p.T = T; p.H = H; p.Hk = Hk; p.Ho = Hk;
  p.qi_st = qi_st; p.qi_sh = qi_sh; p.ki_st = ki_st; p.ki_sh = ki_sh; p.gi_st = gi_st; p.gi_sh = gi_sh;
  p.qo_st = qo_st; p.qo_sh = qo_sh; p.ko_st = ko_st; p.ko_sh = ko_sh; p.go_st = go_st; p.go_sh = go_sh;
  p.clamp_max = 0.f; p.norm = normalizer;
  return launch_gate(p, 2, (cudaStream_t)stream);
}

extern "C" int hymeta_gla_gate_7b_g(const void* gk_in, void* g_out, int T, int Hk, int D, int64_t gi_st,
                                    int64_t gi_sh, int64_t go_st, int64_t go_sh, float normalizer,
                                    void* stream) {
  if (T == 0) return HYMETA_OK;
  if (!gk_in || !g_out || T < 0 || Hk <= 0 || normalizer == 0.f) return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  if ((gi_st % 8) || (gi_sh % 8) || (go_st % 8) || (go_sh % 8)) return HYMETA_ERR_INVALID_ARG;
  GateParams p{};
  p.gk_in = (const __nv_bfloat16*)gk_in; p.g_out = g_out;
  p.T = T; p.H = 0; p.Hk = Hk; p.Ho = Hk;
  p.gi_st = gi_st; p.gi_sh = gi_sh; p.go_st = go_st; p.go_sh = go_sh; p.norm = normalizer;
  return launch_gate(p, 3, (cudaStream_t)stream);
}
