// Branch merge of the 70B hybrid layer (f1): HymetaRMSNormTP of the SWA branch and of the
// GLA branch, their average and the bf16 cast -- modeling_hymeta.py:151-173 (norm) and :909-919
// (norm1(out_attn), norm2(out_linear), (a + b) / 2, .to(bfloat16)) -- in one pass over the two
// [T, hidden/tp] activations instead of ~10 elementwise torch launches with fp32 intermediates.
//
// The norm's variance spans the FULL hidden size, i.e. all tensor-parallel shards:
//   variance = all_reduce_sum( mean_local(x^2) ) / tp_world                         (:159-161)
// so the work is split exactly where the exchange sits:
//   hymeta_branch_sumsq        -> means[2][T] fp32: mean over the local hidden slice of x^2, both branches
//   (caller: ONE all-reduce of 8*T bytes over NVLink -- torch.distributed / NCCL; skipped at TP=1)
//   hymeta_branch_norm_merge   -> out = bf16( ( bf16(x1*rstd1)*w1 + bf16(x2*rstd2)*w2 ) / 2 )
// With means == null the second kernel computes the local means itself (TP=1: one launch).
// Arithmetic is the reference's: x in fp32, rsqrt(variance + eps) in fp32, the normalised value cast
// to bf16 BEFORE the fp32 weight multiplies it (x.to(orig_dtype) * weight promotes to fp32), the
// average in fp32, one final bf16 rounding.  HBM-bound: 2 reads + 1 write of [T, hidden/tp] bf16.
#include "common.cuh"

namespace hymeta {

// One warp per token row; lanes stride over 16-byte chunks, four independent loads in flight per
// lane and branch (a 3584-wide row is 14 chunks per lane: without batching the loop is one
// dependent 16-byte load at a time and the kernel sits at ~55 % of HBM).
constexpr int BM_B = 4;

__device__ __forceinline__ float sumsq8(const uint4& u, float s) {
  const uint32_t* p = &u.x;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float a = bf16_lo(p[e]), b = bf16_hi(p[e]);
    s = fmaf(a, a, s);
    s = fmaf(b, b, s);
  }
  return s;
}

__device__ __forceinline__ void row_sumsq2(const __nv_bfloat16* r1, const __nv_bfloat16* r2, int chunks,
                                           int lane, float& s1, float& s2) {
  s1 = 0.f; s2 = 0.f;
  for (int c0 = lane; c0 < chunks; c0 += 32 * BM_B) {
    uint4 a[BM_B], b[BM_B];
#pragma unroll
    for (int i = 0; i < BM_B; ++i) {
      const int c = c0 + 32 * i;
      a[i] = b[i] = make_uint4(0, 0, 0, 0);
      if (c < chunks) { a[i] = ldg_v4(r1 + c * 8); b[i] = ldg_v4(r2 + c * 8); }
    }
#pragma unroll
    for (int i = 0; i < BM_B; ++i) { s1 = sumsq8(a[i], s1); s2 = sumsq8(b[i], s2); }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, off);
    s2 += __shfl_xor_sync(0xffffffffu, s2, off);
  }
}

__global__ void __launch_bounds__(256)
branch_sumsq_kernel(const __nv_bfloat16* __restrict__ x1, const __nv_bfloat16* __restrict__ x2,
                    float* __restrict__ means, int T, int hidden, int64_t x1_st, int64_t x2_st) {
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const float inv = 1.f / (float)hidden;
  for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < T; t += gridDim.x * wpb) {
    float s1, s2;
    row_sumsq2(x1 + t * x1_st, x2 + t * x2_st, hidden / 8, lane, s1, s2);
    if (lane == 0) { means[t] = s1 * inv; means[T + t] = s2 * inv; }
  }
}

__global__ void __launch_bounds__(256)
branch_norm_merge_kernel(const __nv_bfloat16* __restrict__ x1, const __nv_bfloat16* __restrict__ x2,
                         const float* __restrict__ means, const float* __restrict__ w1,
                         const float* __restrict__ w2, __nv_bfloat16* __restrict__ out, int T,
                         int hidden, float inv_tp, float eps, int64_t x1_st, int64_t x2_st,
                         int64_t o_st) {
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int chunks = hidden / 8;
  const float inv = 1.f / (float)hidden;
  for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < T; t += gridDim.x * wpb) {
    const __nv_bfloat16* r1 = x1 + t * x1_st;
    const __nv_bfloat16* r2 = x2 + t * x2_st;
    float m1, m2;
    if (means) { m1 = means[t]; m2 = means[T + t]; }
    else { row_sumsq2(r1, r2, chunks, lane, m1, m2); m1 *= inv; m2 *= inv; }
    const float rs1 = rsqrtf(m1 * inv_tp + eps), rs2 = rsqrtf(m2 * inv_tp + eps);
    __nv_bfloat16* o = out + t * o_st;
    for (int c0 = lane; c0 < chunks; c0 += 32 * BM_B) {
      uint4 av[BM_B], bv[BM_B];
#pragma unroll
      for (int i = 0; i < BM_B; ++i) {                   // second touch of the row: L1/L2 hits
        const int c = c0 + 32 * i;
        if (c < chunks) { av[i] = ldg_v4(r1 + c * 8); bv[i] = ldg_v4(r2 + c * 8); }
      }
#pragma unroll
      for (int i = 0; i < BM_B; ++i) {
        const int c = c0 + 32 * i;
        if (c >= chunks) break;
        const float4 wa0 = *reinterpret_cast<const float4*>(w1 + c * 8);
        const float4 wa1 = *reinterpret_cast<const float4*>(w1 + c * 8 + 4);
        const float4 wb0 = *reinterpret_cast<const float4*>(w2 + c * 8);
        const float4 wb1 = *reinterpret_cast<const float4*>(w2 + c * 8 + 4);
        const float wa[8] = {wa0.x, wa0.y, wa0.z, wa0.w, wa1.x, wa1.y, wa1.z, wa1.w};
        const float wb[8] = {wb0.x, wb0.y, wb0.z, wb0.w, wb1.x, wb1.y, wb1.z, wb1.w};
        const uint32_t* ap = &av[i].x;
        const uint32_t* bp = &bv[i].x;
        uint4 r;
        uint32_t* rp = &r.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float y0 = __fadd_rn(__fmul_rn(round_bf16(__fmul_rn(bf16_lo(ap[e]), rs1)), wa[2 * e]),
                                     __fmul_rn(round_bf16(__fmul_rn(bf16_lo(bp[e]), rs2)), wb[2 * e]));
          const float y1 = __fadd_rn(__fmul_rn(round_bf16(__fmul_rn(bf16_hi(ap[e]), rs1)), wa[2 * e + 1]),
                                     __fmul_rn(round_bf16(__fmul_rn(bf16_hi(bp[e]), rs2)), wb[2 * e + 1]));
          rp[e] = pack_bf16(y0 * 0.5f, y1 * 0.5f);
        }
        stg_v4(o + c * 8, r);
      }
    }
  }
}

// Register-resident form for rows of at most 32 * NC chunks (NC = 14: hidden/tp <= 3584, every
// Hymeta geometry): the warp loads both rows ONCE into registers (2 * NC independent 16-byte loads
// in flight per lane), reduces, scales and stores -- no second touch, no dependent load chain.
template <int NC>
__global__ void __launch_bounds__(128)
branch_norm_merge_reg_kernel(const __nv_bfloat16* x1, const __nv_bfloat16* x2, const float* means,
                             const float* __restrict__ w1, const float* __restrict__ w2,
                             __nv_bfloat16* out, int T, int hidden, float inv_tp, float eps,
                             int64_t x1_st, int64_t x2_st, int64_t o_st) {
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  const int chunks = hidden / 8;
  const float inv = 1.f / (float)hidden;
  for (int t = blockIdx.x * wpb + (threadIdx.x >> 5); t < T; t += gridDim.x * wpb) {
    const uint4* r1 = reinterpret_cast<const uint4*>(x1 + t * x1_st);
    const uint4* r2 = reinterpret_cast<const uint4*>(x2 + t * x2_st);
    uint4 a[NC], b[NC];
#pragma unroll
    for (int i = 0; i < NC; ++i) {
      const int c = lane + 32 * i;
      a[i] = b[i] = make_uint4(0, 0, 0, 0);
      if (c < chunks) { a[i] = r1[c]; b[i] = r2[c]; }
    }
    float m1, m2;
    if (means) { m1 = means[t]; m2 = means[T + t]; }
    else {
      float s1 = 0.f, s2 = 0.f;
#pragma unroll
      for (int i = 0; i < NC; ++i) { s1 = sumsq8(a[i], s1); s2 = sumsq8(b[i], s2); }
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        s1 += __shfl_xor_sync(0xffffffffu, s1, off);
        s2 += __shfl_xor_sync(0xffffffffu, s2, off);
      }
      m1 = s1 * inv; m2 = s2 * inv;
    }
    const float rs1 = rsqrtf(m1 * inv_tp + eps), rs2 = rsqrtf(m2 * inv_tp + eps);
    uint4* o = reinterpret_cast<uint4*>(out + t * o_st);
#pragma unroll
    for (int i = 0; i < NC; ++i) {
      const int c = lane + 32 * i;
      if (c < chunks) {
        const float4 wa0 = *reinterpret_cast<const float4*>(w1 + c * 8);
        const float4 wa1 = *reinterpret_cast<const float4*>(w1 + c * 8 + 4);
        const float4 wb0 = *reinterpret_cast<const float4*>(w2 + c * 8);
        const float4 wb1 = *reinterpret_cast<const float4*>(w2 + c * 8 + 4);
        const float wa[8] = {wa0.x, wa0.y, wa0.z, wa0.w, wa1.x, wa1.y, wa1.z, wa1.w};
        const float wb[8] = {wb0.x, wb0.y, wb0.z, wb0.w, wb1.x, wb1.y, wb1.z, wb1.w};
        const uint32_t* ap = &a[i].x;
        const uint32_t* bp = &b[i].x;
        uint4 r;
        uint32_t* rp = &r.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float y0 = __fadd_rn(__fmul_rn(round_bf16(__fmul_rn(bf16_lo(ap[e]), rs1)), wa[2 * e]),
                                     __fmul_rn(round_bf16(__fmul_rn(bf16_lo(bp[e]), rs2)), wb[2 * e]));
          const float y1 = __fadd_rn(__fmul_rn(round_bf16(__fmul_rn(bf16_hi(ap[e]), rs1)), wa[2 * e + 1]),
                                     __fmul_rn(round_bf16(__fmul_rn(bf16_hi(bp[e]), rs2)), wb[2 * e + 1]));
          rp[e] = pack_bf16(y0 * 0.5f, y1 * 0.5f);
        }
        o[c] = r;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Tensor-parallel form, ONE kernel: local mean of squares -> exchange over NVLink peer memory ->
// normalise + merge (x is read from DRAM once: the second touch of a row, a few microseconds after
// the first, is an L2 hit).
//
// Exchange = the low-latency flag-in-data protocol: every rank owns a peer-mapped buffer
//   uint4 slots[2 parities][world][t_cap]     slot = {mean_attn bits, epoch, mean_linear bits, epoch}
// and a warp that has reduced row t writes its slot [epoch & 1][rank][t] INTO EVERY PEER'S buffer with
// one 16-byte store (each 8-byte half carries its own epoch, so a torn store is still detected), then
// polls the peers' slots for row t in its OWN buffer until their epochs match.  No fence, no separate
// flag, no grid-wide synchronisation: rows are independent, and the warps in flight on an SM hide the
// ~3 us NVLink round trip.  The sum over ranks is taken in rank order on every rank (bit-identical
// variance everywhere).  Two parities: a peer may start call n+1 (and overwrite its slots) while this
// rank still reads call n, but cannot reach call n+2 before this rank has finished call n+1.
struct TpPeers { uint4* buf[8]; };

constexpr int TP_K = 4;   // rows per warp between "push" and "poll": the exchange latency of a row is
                          // covered by the other rows' loads; the second touch of the rows hits L2

__global__ void __launch_bounds__(128)
branch_norm_merge_tp_kernel(const __nv_bfloat16* x1, const __nv_bfloat16* x2,
                            const float* __restrict__ w1, const float* __restrict__ w2,
                            __nv_bfloat16* out, int T, int hidden, int world, int rank, uint32_t epoch,
                            int64_t t_cap, const TpPeers peers, float eps, int64_t x1_st, int64_t x2_st,
                            int64_t o_st, const float* __restrict__ ss1, const float* __restrict__ ss2,
                            int heads, uint32_t* epoch_state) {
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
  if (epoch_state) {
    // Device-side epoch (CUDA-graph replays carry no fresh by-value argument): {last epoch, CTAs done}.
    // Every CTA of this launch reads the same last epoch -- it is advanced by the LAST CTA to finish.
    uint32_t cur;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(cur) : "l"(epoch_state) : "memory");
    epoch = cur % 0x7fffffffu + 1u;              // never 0 (the zero fill of the slots)
  }
  const int chunks = hidden / 8;
  const float inv = 1.f / (float)hidden, inv_tp = 1.f / (float)world;
  const int64_t par = (int64_t)(epoch & 1u) * world;
  // lane r < world talks to rank r: where my slot lives in r's buffer, where r's slot lives in mine
  uint4* push = nullptr;
  const uint4* poll = nullptr;
  if (lane < world && lane != rank) {
    push = peers.buf[lane] + (par + rank) * t_cap;
    poll = peers.buf[rank] + (par + lane) * t_cap;
  }
  const int stride = gridDim.x * wpb;
  for (int t0 = blockIdx.x * wpb + (threadIdx.x >> 5); t0 < T; t0 += stride * TP_K) {
    float m1[TP_K], m2[TP_K];
    // (A) local means of TP_K rows, each pushed to every peer as soon as it is reduced
#pragma unroll
    for (int j = 0; j < TP_K; ++j) {
      const int t = t0 + j * stride;
      m1[j] = m2[j] = 0.f;
      if (t < T) {
        float s1, s2;
        if (ss1) {   // per-head partials from the attention kernels' epilogues: x is not read here
          s1 = s2 = 0.f;
          for (int h = lane; h < heads; h += 32) { s1 += ss1[(int64_t)t * heads + h]; s2 += ss2[(int64_t)t * heads + h]; }
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) {
            s1 += __shfl_xor_sync(0xffffffffu, s1, off);
            s2 += __shfl_xor_sync(0xffffffffu, s2, off);
          }
        } else {
          row_sumsq2(x1 + t * x1_st, x2 + t * x2_st, chunks, lane, s1, s2);
        }
        m1[j] = s1 * inv; m2[j] = s2 * inv;
        if (push)
          asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};"
                       :: "l"(push + t), "r"(__float_as_uint(m1[j])), "r"(epoch), "r"(__float_as_uint(m2[j])),
                          "r"(epoch) : "memory");
      }
    }
    // (B) the peers' means of the same rows (lane r polls rank r's slots in this rank's buffer)
#pragma unroll
    for (int j = 0; j < TP_K; ++j) {
      const int t = t0 + j * stride;
      if (t < T) {
        float a = m1[j], b = m2[j];
        if (poll) {
          uint32_t x, ex, y, ey, spins = 0;
          do {
            asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(x), "=r"(ex), "=r"(y), "=r"(ey) : "l"(poll + t) : "memory");
            // a peer that never arrives (a rank died, or the ranks disagree on the call sequence) must
            // not hang the GPU: give up after ~1-2 minutes of polling -- far beyond any legitimate skew
            // between ranks -- and fail the launch loudly
            if (++spins > (1u << 27)) asm volatile("trap;");
          } while (ex != epoch || ey != epoch);
          a = __uint_as_float(x); b = __uint_as_float(y);
        }
        float t1 = 0.f, t2 = 0.f;
        for (int r = 0; r < world; ++r) {                        // rank order: the same sum on every rank
          t1 += __shfl_sync(0xffffffffu, a, r);
          t2 += __shfl_sync(0xffffffffu, b, r);
        }
        m1[j] = rsqrtf(t1 * inv_tp + eps); m2[j] = rsqrtf(t2 * inv_tp + eps);
      }
    }
    // (C) normalise + merge: the rows were read a few microseconds ago and come from L2
#pragma unroll
    for (int j = 0; j < TP_K; ++j) {
      const int t = t0 + j * stride;
      if (t >= T) break;
      const __nv_bfloat16* r1 = x1 + t * x1_st;
      const __nv_bfloat16* r2 = x2 + t * x2_st;
      const float rs1 = m1[j], rs2 = m2[j];
      __nv_bfloat16* o = out + t * o_st;
      for (int c0 = lane; c0 < chunks; c0 += 32 * BM_B) {
        uint4 av[BM_B], bv[BM_B];
#pragma unroll
        for (int i = 0; i < BM_B; ++i) {
          const int c = c0 + 32 * i;
          if (c < chunks) { av[i] = ldg_v4(r1 + c * 8); bv[i] = ldg_v4(r2 + c * 8); }
        }
#pragma unroll
        for (int i = 0; i < BM_B; ++i) {
          const int c = c0 + 32 * i;
          if (c >= chunks) break;
          const float4 wa0 = *reinterpret_cast<const float4*>(w1 + c * 8);
          const float4 wa1 = *reinterpret_cast<const float4*>(w1 + c * 8 + 4);
          const float4 wb0 = *reinterpret_cast<const float4*>(w2 + c * 8);
          const float4 wb1 = *reinterpret_cast<const float4*>(w2 + c * 8 + 4);
          const float wa[8] = {wa0.x, wa0.y, wa0.z, wa0.w, wa1.x, wa1.y, wa1.z, wa1.w};
          const float wb[8] = {wb0.x, wb0.y, wb0.z, wb0.w, wb1.x, wb1.y, wb1.z, wb1.w};
          const uint32_t* ap = &av[i].x;
          const uint32_t* bp = &bv[i].x;
          uint4 r;
          uint32_t* rp = &r.x;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float y0 = __fadd_rn(__fmul_rn(round_bf16(__fmul_rn(bf16_lo(ap[e]), rs1)), wa[2 * e]),
                                       __fmul_rn(round_bf16(__fmul_rn(bf16_lo(bp[e]), rs2)), wb[2 * e]));
            const float y1 = __fadd_rn(__fmul_rn(round_bf16(__fmul_rn(bf16_hi(ap[e]), rs1)), wa[2 * e + 1]),
                                       __fmul_rn(round_bf16(__fmul_rn(bf16_hi(bp[e]), rs2)), wb[2 * e + 1]));
            rp[e] = pack_bf16(y0 * 0.5f, y1 * 0.5f);
          }
          stg_v4(o + c * 8, r);
        }
      }
    }
  }
  if (epoch_state) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(epoch_state + 1, 1u) == gridDim.x - 1) {
        epoch_state[1] = 0;
        asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(epoch_state), "r"(epoch) : "memory");
      }
    }
  }
}

static int check_common(const void* x1, const void* x2, int T, int hidden, int64_t x1_st,
                        int64_t x2_st) {
  if (!x1 || !x2 || T < 0 || hidden <= 0) return HYMETA_ERR_INVALID_ARG;
  if (hidden % 8 || x1_st % 8 || x2_st % 8) return HYMETA_ERR_INVALID_ARG;   // 16-byte vector access
  if (((uintptr_t)x1 | (uintptr_t)x2) & 15) return HYMETA_ERR_INVALID_ARG;
  return HYMETA_OK;
}
static int grid_for(int T) {
  const int want = (T + 7) / 8;   // 8 warps (rows) per CTA
  return want < 148 * 8 ? want : 148 * 8;
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_branch_sumsq(const void* x_attn, const void* x_linear, float* means_out, int T,
                                   int hidden_local, int64_t xa_st, int64_t xl_st, void* stream) {
  if (T == 0) return HYMETA_OK;
  const int rc = check_common(x_attn, x_linear, T, hidden_local, xa_st, xl_st);
  if (rc) return rc;
  if (!means_out) return HYMETA_ERR_INVALID_ARG;
  branch_sumsq_kernel<<<grid_for(T), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)x_attn, (const __nv_bfloat16*)x_linear, means_out, T, hidden_local,
      xa_st, xl_st);
  count_launch();
  return check_launch();
}

extern "C" int hymeta_branch_norm_merge(const void* x_attn, const void* x_linear, const float* means,
                                        const float* w_attn, const float* w_linear, void* out, int T,
                                        int hidden_local, int tp_world, float eps, int64_t xa_st,
                                        int64_t xl_st, int64_t o_st, void* stream) {
  if (T == 0) return HYMETA_OK;
  const int rc = check_common(x_attn, x_linear, T, hidden_local, xa_st, xl_st);
  if (rc) return rc;
  if (!w_attn || !w_linear || !out || tp_world <= 0 || o_st % 8) return HYMETA_ERR_INVALID_ARG;
  if (((uintptr_t)out | (uintptr_t)w_attn | (uintptr_t)w_linear) & 15) return HYMETA_ERR_INVALID_ARG;
  if (tp_world > 1 && !means) return HYMETA_ERR_INVALID_ARG;   // the cross-shard sum is the caller's
  const float inv_tp = 1.f / (float)tp_world;
  if (hidden_local <= 32 * 8 * 14) {   // row fits the register-resident form
    const int want = (T + 3) / 4;      // 4 warps (rows) per CTA
    const int grid = want < 148 * 4 ? want : 148 * 4;
    if (hidden_local <= 32 * 8 * 4)
      branch_norm_merge_reg_kernel<4><<<grid, 128, 0, (cudaStream_t)stream>>>(
          (const __nv_bfloat16*)x_attn, (const __nv_bfloat16*)x_linear, means, w_attn, w_linear,
          (__nv_bfloat16*)out, T, hidden_local, inv_tp, eps, xa_st, xl_st, o_st);
    else
      branch_norm_merge_reg_kernel<14><<<grid, 128, 0, (cudaStream_t)stream>>>(
          (const __nv_bfloat16*)x_attn, (const __nv_bfloat16*)x_linear, means, w_attn, w_linear,
          (__nv_bfloat16*)out, T, hidden_local, inv_tp, eps, xa_st, xl_st, o_st);
  } else {
    branch_norm_merge_kernel<<<grid_for(T), 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x_attn, (const __nv_bfloat16*)x_linear, means, w_attn, w_linear,
        (__nv_bfloat16*)out, T, hidden_local, inv_tp, eps, xa_st, xl_st, o_st);
  }
  count_launch();
  return check_launch();
}

extern "C" size_t hymeta_branch_merge_tp_buffer_bytes(int tp_world, int64_t t_cap) {
  if (tp_world <= 1 || tp_world > 8 || t_cap <= 0) return 0;
  return (size_t)2 * tp_world * (size_t)t_cap * sizeof(uint4);
}

static int branch_norm_merge_tp_impl(const void* x_attn, const void* x_linear, const float* w_attn,
                                     const float* w_linear, void* out, int T, int hidden_local,
                                     int tp_world, int tp_rank, const uint64_t* peer_buffers,
                                     int64_t t_cap, uint32_t epoch, uint32_t* epoch_state, float eps,
                                     int64_t xa_st, int64_t xl_st, int64_t o_st, const float* sumsq_attn,
                                     const float* sumsq_linear, int heads_local, void* stream) {
  // T == 0 launches nothing on any rank (every rank passes the same T), so the epochs stay aligned
  if (T == 0) return HYMETA_OK;
  if (epoch_state) {
    if ((uintptr_t)epoch_state & 7) return HYMETA_ERR_INVALID_ARG;
    epoch = 1;   // placeholder: the kernel derives the epoch from epoch_state
  }
  if ((sumsq_attn == nullptr) != (sumsq_linear == nullptr) || (sumsq_attn && heads_local <= 0))
    return HYMETA_ERR_INVALID_ARG;
  const int rc = check_common(x_attn, x_linear, T, hidden_local, xa_st, xl_st);
  if (rc) return rc;
  if (!w_attn || !w_linear || !out || !peer_buffers || o_st % 8) return HYMETA_ERR_INVALID_ARG;
  if (tp_world < 2 || tp_world > 8 || tp_rank < 0 || tp_rank >= tp_world || T > t_cap || epoch == 0)
    return HYMETA_ERR_INVALID_ARG;
  if (((uintptr_t)out | (uintptr_t)w_attn | (uintptr_t)w_linear) & 15) return HYMETA_ERR_INVALID_ARG;
  TpPeers peers{};
  for (int r = 0; r < tp_world; ++r) {
    if (!peer_buffers[r] || (peer_buffers[r] & 15)) return HYMETA_ERR_INVALID_ARG;
    peers.buf[r] = reinterpret_cast<uint4*>(peer_buffers[r]);
  }
  const int want = (T + 3) / 4;                  // 4 warps (rows) per CTA
  auto go = [&](auto kernel) {
    // every CTA of the grid must be resident at once (a warp polls for rows a peer's CTA of the same
    // index produces): grid <= SMs x occupancy; the rows in flight per SM hide the exchange latency
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, 128, 0) != cudaSuccess || occ < 1) occ = 1;
    const int cap = sm_count() * occ;
    const int grid = want < cap ? want : cap;
    kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x_attn, (const __nv_bfloat16*)x_linear, w_attn, w_linear, (__nv_bfloat16*)out, T,
        hidden_local, tp_world, tp_rank, epoch, t_cap, peers, eps, xa_st, xl_st, o_st, sumsq_attn, sumsq_linear,
        heads_local, epoch_state);
  };
  go(branch_norm_merge_tp_kernel);
  count_launch();
  return check_launch();
}

extern "C" int hymeta_branch_norm_merge_tp_sumsq(const void* x_attn, const void* x_linear, const float* w_attn,
                                           const float* w_linear, void* out, int T, int hidden_local,
                                           int tp_world, int tp_rank, const uint64_t* peer_buffers,
                                           int64_t t_cap, uint32_t epoch, float eps, int64_t xa_st,
                                           int64_t xl_st, int64_t o_st, const float* sumsq_attn,
                                           const float* sumsq_linear, int heads_local, void* stream) {
  return branch_norm_merge_tp_impl(x_attn, x_linear, w_attn, w_linear, out, T, hidden_local, tp_world, tp_rank,
                                   peer_buffers, t_cap, epoch, nullptr, eps, xa_st, xl_st, o_st, sumsq_attn,
                                   sumsq_linear, heads_local, stream);
}

extern "C" int hymeta_branch_norm_merge_tp_graph(const void* x_attn, const void* x_linear, const float* w_attn,
                                           const float* w_linear, void* out, int T, int hidden_local,
                                           int tp_world, int tp_rank, const uint64_t* peer_buffers,
                                           int64_t t_cap, uint32_t* epoch_state, float eps, int64_t xa_st,
                                           int64_t xl_st, int64_t o_st, const float* sumsq_attn,
                                           const float* sumsq_linear, int heads_local, void* stream) {
  if (!epoch_state) return HYMETA_ERR_INVALID_ARG;
  return branch_norm_merge_tp_impl(x_attn, x_linear, w_attn, w_linear, out, T, hidden_local, tp_world, tp_rank,
                                   peer_buffers, t_cap, 0, epoch_state, eps, xa_st, xl_st, o_st, sumsq_attn,
                                   sumsq_linear, heads_local, stream);
}

extern "C" int hymeta_branch_norm_merge_tp(const void* x_attn, const void* x_linear, const float* w_attn,
                                           const float* w_linear, void* out, int T, int hidden_local,
                                           int tp_world, int tp_rank, const uint64_t* peer_buffers,
                                           int64_t t_cap, uint32_t epoch, float eps, int64_t xa_st,
                                           int64_t xl_st, int64_t o_st, void* stream) {
  return hymeta_branch_norm_merge_tp_sumsq(x_attn, x_linear, w_attn, w_linear, out, T, hidden_local, tp_world,
                                           tp_rank, peer_buffers, t_cap, epoch, eps, xa_st, xl_st, o_st, nullptr,
                                           nullptr, 0, stream);
}
