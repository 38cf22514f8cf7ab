// GLA recurrent decode step (a1/a2) -- HBM-bound read-modify-write of the
// per-(request, state-head) 128x128 bf16 state.
//
//   S <- bf16(k (x) v) + exp(g)[:,None] * S        (fp32 math, bf16 store)
//   o  = (q * scale)^T S                            (all q-heads sharing the state)
//
// Grid (D/DVS, Hs, B): one CTA owns all 128 d_k rows of a DVS-wide d_v slice, so the reduction
// over d_k stays inside the CTA (warp shuffles + one smem pass) and every state byte is touched
// by exactly one thread with 128-bit loads/stores.  The G = H/Hs q-heads of a GQA group are all
// served from the single state read.
#include "common.cuh"

namespace hymeta {

struct GlaDecodeParams {
  const __nv_bfloat16 *q, *k, *v;
  const void* g;
  __nv_bfloat16* o;
  __nv_bfloat16* state;
  const int32_t* slot_idx;
  int B, H, Hk, Hs;
  int64_t q_sb, q_sh, k_sb, k_sh, v_sb, v_sh, g_sb, g_sh, o_sb, o_sh, st_ss, st_sh;
  float scale;
};

template <int DVS, int GT, bool GF32>
__global__ void __launch_bounds__(256) gla_decode_kernel(const GlaDecodeParams p) {
  constexpr int TPR = DVS / 8;    // threads per state row (16 B each)
  constexpr int RPP = 256 / TPR;  // rows per pass
  constexpr int NP = kD / RPP;    // passes
  const int b = blockIdx.z, hs = blockIdx.y, dv0 = blockIdx.x * DVS;
  const int G = p.H / p.Hs;
  const int hk = hs / (p.Hs / p.Hk);
  const int h0 = hs * G;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int c = tid % TPR, r0 = tid / TPR;

  const int slot = p.slot_idx[b];
  if (slot < 0) {  // PAD slot (CUDA-graph padding): skip, emit zeros
    for (int i = tid; i < G * DVS; i += 256)
      p.o[b * p.o_sb + (h0 + i / DVS) * p.o_sh + dv0 + (i % DVS)] = __float2bfloat16(0.f);
    return;
  }

  __shared__ float s_k[kD], s_eg[kD];
  __shared__ float s_q[GT][kD];
  __shared__ float s_red[8][GT][DVS];

  // 1. state loads first: they head the longest dependency chain
  __nv_bfloat16* sp = p.state + (int64_t)slot * p.st_ss + (int64_t)hs * p.st_sh + dv0 + c * 8;
  uint4 sv[NP];
#pragma unroll
  for (int i = 0; i < NP; ++i) sv[i] = ldg_v4(sp + (i * RPP + r0) * kD);

  // 2. k, exp(g) -> smem; v slice -> registers
  if (tid < kD) {
    s_k[tid] = __bfloat162float(p.k[b * p.k_sb + hk * p.k_sh + tid]);
    float gv;
    if (GF32) gv = reinterpret_cast<const float*>(p.g)[b * p.g_sb + hk * p.g_sh + tid];
    else gv = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.g)[b * p.g_sb + hk * p.g_sh + tid]);
    s_eg[tid] = expf(gv);
  }
  float vv[8];
  {
    uint4 t = ldg_nc_v4(p.v + b * p.v_sb + hk * p.v_sh + dv0 + c * 8);
    vv[0] = bf16_lo(t.x); vv[1] = bf16_hi(t.x); vv[2] = bf16_lo(t.y); vv[3] = bf16_hi(t.y);
    vv[4] = bf16_lo(t.z); vv[5] = bf16_hi(t.z); vv[6] = bf16_lo(t.w); vv[7] = bf16_hi(t.w);
  }

  float sn[NP][8];  // new state (fp32, un-rounded: the output reduces this, my_fused_recurrent.py:75)
  for (int hc = 0; hc < G; hc += GT) {
    const int ng = min(GT, G - hc);
    for (int i = tid; i < ng * kD; i += 256)
      s_q[i / kD][i % kD] =
          __bfloat162float(p.q[b * p.q_sb + (h0 + hc + i / kD) * p.q_sh + (i % kD)]) * p.scale;
    __syncthreads();

    float acc[GT][8];
#pragma unroll
    for (int hh = 0; hh < GT; ++hh)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[hh][j] = 0.f;

#pragma unroll
    for (int i = 0; i < NP; ++i) {
      const int r = i * RPP + r0;
      if (hc == 0) {
        const float kr = s_k[r], eg = s_eg[r];
        const uint32_t w[4] = {sv[i].x, sv[i].y, sv[i].z, sv[i].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          sn[i][2 * j] = round_bf16(kr * vv[2 * j]) + eg * bf16_lo(w[j]);
          sn[i][2 * j + 1] = round_bf16(kr * vv[2 * j + 1]) + eg * bf16_hi(w[j]);
        }
        uint4 st;
        st.x = pack_bf16(sn[i][0], sn[i][1]); st.y = pack_bf16(sn[i][2], sn[i][3]);
        st.z = pack_bf16(sn[i][4], sn[i][5]); st.w = pack_bf16(sn[i][6], sn[i][7]);
        stg_v4(sp + r * kD, st);
      }
#pragma unroll
      for (int hh = 0; hh < GT; ++hh) {
        if (hh < ng) {
          const float qr = s_q[hh][r];
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[hh][j] = fmaf(qr, sn[i][j], acc[hh][j]);
        }
      }
    }

    // reduce over d_k: lanes sharing the column chunk, then the 8 warps
#pragma unroll
    for (int hh = 0; hh < GT; ++hh) {
      if (hh < ng) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float x = acc[hh][j];
#pragma unroll
          for (int off = 16; off >= TPR; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
          acc[hh][j] = x;
        }
        if (lane < TPR) {
#pragma unroll
          for (int j = 0; j < 8; ++j) s_red[warp][hh][c * 8 + j] = acc[hh][j];
        }
      }
    }
    __syncthreads();
    for (int i = tid; i < ng * DVS; i += 256) {
      const int hh = i / DVS, j = i % DVS;
      float x = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) x += s_red[w][hh][j];
      p.o[b * p.o_sb + (h0 + hc + hh) * p.o_sh + dv0 + j] = __float2bfloat16_rn(x);
    }
    __syncthreads();
  }
}

template <int DVS, int GT>
static void launch_gla_decode(const GlaDecodeParams& p, bool g_fp32, cudaStream_t st) {
  dim3 grid(kD / DVS, p.Hs, p.B), block(256);
  if (g_fp32) gla_decode_kernel<DVS, GT, true><<<grid, block, 0, st>>>(p);
  else gla_decode_kernel<DVS, GT, false><<<grid, block, 0, st>>>(p);
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_gla_decode(const void* q, const void* k, const void* v, const void* g, void* o,
                                 void* state, const int32_t* slot_idx, int B, int H, int Hk, int Hs,
                                 int D, int64_t q_sb, int64_t q_sh, int64_t k_sb, int64_t k_sh,
                                 int64_t v_sb, int64_t v_sh, int64_t g_sb, int64_t g_sh,
                                 int64_t o_sb, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                                 int g_is_fp32, float scale, void* stream) {
  if (B == 0) return HYMETA_OK;
  if (!q || !k || !v || !g || !o || !state || !slot_idx || B < 0 || H <= 0 || Hk <= 0 || Hs <= 0)
    return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  if (H % Hs != 0 || Hs % Hk != 0) return HYMETA_ERR_INVALID_ARG;
  if ((st_ss % 8) || (st_sh % 8) || (v_sb % 8) || (v_sh % 8)) return HYMETA_ERR_INVALID_ARG;
  if (B > 65535) return HYMETA_ERR_UNSUPPORTED;
  GlaDecodeParams p{(const __nv_bfloat16*)q, (const __nv_bfloat16*)k, (const __nv_bfloat16*)v, g,
                    (__nv_bfloat16*)o, (__nv_bfloat16*)state, slot_idx, B, H, Hk, Hs,
                    q_sb, q_sh, k_sb, k_sh, v_sb, v_sh, g_sb, g_sh, o_sb, o_sh, st_ss, st_sh, scale};
  cudaStream_t st = (cudaStream_t)stream;
  const int G = H / Hs;
  const long states = (long)B * Hs;
  // enough CTAs to cover 148 SMs a few times over; wider slices once the batch alone does that
  const int dvs = states >= 1184 ? 128 : (states >= 592 ? 64 : 32);
  const bool f = g_is_fp32 != 0;
  if (G == 1) {
    if (dvs == 128) launch_gla_decode<128, 1>(p, f, st);
    else if (dvs == 64) launch_gla_decode<64, 1>(p, f, st);
    else launch_gla_decode<32, 1>(p, f, st);
  } else {
    if (dvs == 128) launch_gla_decode<128, 8>(p, f, st);
    else if (dvs == 64) launch_gla_decode<64, 8>(p, f, st);
    else launch_gla_decode<32, 8>(p, f, st);
  }
  count_launch();
  return check_launch();
}
