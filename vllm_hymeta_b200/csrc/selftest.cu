// Self test of the tcgen05 / TMEM / TMA building blocks: every shared-memory descriptor flavour the
// prefill kernels rely on (K-major and MN-major A and B, SWIZZLE_128B, generic-proxy writes and TMA
// writes, N = 16/64/128, K = 16/64/128) against a CUDA-core GEMM.
#include <stdio.h>

#include <vector>

#include "umma.cuh"

namespace hymeta {

struct SelfCase { int a_mn, b_mn, N, K, tma, a_tmem; };   // a_tmem: A staged in TMEM (tcgen05.st), K-major

// A: !a_mn -> [128][K] (K contiguous), a_mn -> [K][128].  B: !b_mn -> [N][K], b_mn -> [K][N].
__global__ void __launch_bounds__(128)
selftest_umma_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B,
                     float* __restrict__ D, SelfCase c, const __grid_constant__ CUtensorMap tmA,
                     const __grid_constant__ CUtensorMap tmB) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  const uint32_t sA = smem_u32(smem);
  const uint32_t sB = sA + 128 * 128 * 2;  // A occupies <= 32 KB
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t bar_tma = smem_u32(&bars[0]), bar_mma = smem_u32(&bars[1]);
  const int N = c.N, K = c.K;

  if (tid == 0) {
    mbar_init(bar_tma, 1);
    mbar_init(bar_mma, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<256>(smem_u32(&tmem_base_s));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;

  if (c.tma) {
    if (tid == 0) {
      const uint32_t bytes = (128 * K + N * K) * 2;
      mbar_expect_tx(bar_tma, bytes);
      // A
      if (!c.a_mn) for (int kb = 0; kb < K / 64; ++kb) tma_load_3d(sA + kb * 128 * 128, &tmA, bar_tma, kb * 64, 0, 0);
      else for (int mc = 0; mc < 2; ++mc) tma_load_3d(sA + mc * K * 128, &tmA, bar_tma, mc * 64, 0, 0);
      if (!c.b_mn) for (int kb = 0; kb < K / 64; ++kb) tma_load_3d(sB + kb * N * 128, &tmB, bar_tma, kb * 64, 0, 0);
      else for (int nc = 0; nc < N / 64; ++nc) tma_load_3d(sB + nc * K * 128, &tmB, bar_tma, nc * 64, 0, 0);
    }
    mbar_wait(bar_tma, 0);
  } else {
    for (int i = tid; i < 128 * K; i += 128) {
      int m, k;
      if (!c.a_mn) { m = i / K; k = i % K; } else { k = i / 128; m = i % 128; }
      uint32_t off = !c.a_mn ? (k / 64) * 128 * 128 + sw128_off(m, k % 64)
                             : (m / 64) * K * 128 + sw128_off(k, m % 64);
      *reinterpret_cast<__nv_bfloat16*>(smem + off) = A[i];
    }
    for (int i = tid; i < N * K; i += 128) {
      int n, k;
      if (!c.b_mn) { n = i / K; k = i % K; } else { k = i / N; n = i % N; }
      uint32_t off = !c.b_mn ? (k / 64) * N * 128 + sw128_off(n, k % 64)
                             : (n / 64) * K * 128 + sw128_off(k, n % 64);
      *reinterpret_cast<__nv_bfloat16*>(smem + 128 * 128 * 2 + off) = B[i];
    }
    fence_proxy_async_smem();
  }
  if (c.a_tmem) {   // row tid of A -> lane tid, bf16 pairs in columns [128, 128 + K/2)
    for (int c0 = 0; c0 < K / 2; c0 += 16) {
      uint32_t w[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const __nv_bfloat16 lo = A[tid * K + 2 * (c0 + j)], hi = A[tid * K + 2 * (c0 + j) + 1];
        w[j] = (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
      }
      tmem_st16(tmem_base_s + ((uint32_t)(warp * 32) << 16) + 128 + c0, w);
    }
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (tid == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, N, c.a_mn, c.b_mn);
    for (int ks = 0; ks < K / 16; ++ks) {
      uint64_t da = !c.a_mn ? umma_desc_sw128(sA + (ks / 4) * 128 * 128 + (ks % 4) * 32, 16, 1024)
                            : umma_desc_sw128(sA + ks * 2048, K * 128, 1024);
      uint64_t db = !c.b_mn ? umma_desc_sw128(sB + (ks / 4) * N * 128 + (ks % 4) * 32, 16, 1024)
                            : umma_desc_sw128(sB + ks * 2048, K * 128, 1024);
      if (c.a_tmem) umma_bf16_ts(tmem, tmem + 128 + ks * 8, db, idesc, ks > 0);
      else umma_bf16(tmem, da, db, idesc, ks > 0);
    }
    umma_commit(bar_mma);
  }
  mbar_wait(bar_mma, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < N; c0 += 16) {
    float v[16];
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
#pragma unroll
    for (int j = 0; j < 16; ++j) D[tid * N + c0 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<256>(tmem);
}

__global__ void selftest_ref_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D, SelfCase c) {
  const int m = blockIdx.x, n = threadIdx.x;
  if (n >= c.N) return;
  float acc = 0.f;
  for (int k = 0; k < c.K; ++k) {
    float a = __bfloat162float(c.a_mn ? A[k * 128 + m] : A[m * c.K + k]);
    float b = __bfloat162float(c.b_mn ? B[k * c.N + n] : B[n * c.K + k]);
    acc += a * b;
  }
  D[m * c.N + n] = acc;
}

__global__ void selftest_fill_kernel(__nv_bfloat16* p, int n, uint32_t seed) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t x = (uint32_t)i * 2654435761u + seed;
  x ^= x >> 15; x *= 2246822519u; x ^= x >> 13;
  p[i] = __float2bfloat16(((int)(x % 17) - 8) * 0.125f);   // exact in bf16, sums exact in fp32
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_selftest_umma(float* max_err_out) {
  const SelfCase cases[] = {
      {0, 0, 128, 64, 0}, {0, 0, 128, 64, 1}, {0, 0, 64, 128, 1}, {0, 0, 16, 128, 0},
      {0, 1, 128, 64, 0}, {0, 1, 128, 64, 1}, {0, 1, 128, 16, 0}, {0, 1, 128, 128, 1},
      {1, 0, 128, 64, 0}, {1, 1, 128, 16, 0}, {1, 1, 128, 64, 1}, {0, 0, 128, 16, 0},
      {0, 1, 64, 64, 1},
      {0, 1, 128, 128, 1, 1}, {0, 1, 128, 32, 0, 1}, {0, 0, 64, 64, 0, 1},   // A from TMEM (P @ V)
  };
  __nv_bfloat16 *A, *B;
  float *D, *R;
  if (cudaMalloc(&A, 128 * 128 * 2) != cudaSuccess) return HYMETA_ERR_NO_DEVICE;
  cudaMalloc(&B, 128 * 128 * 2);
  cudaMalloc(&D, 128 * 128 * 4);
  cudaMalloc(&R, 128 * 128 * 4);
  const int smem = 64 * 1024 + 1024;
  cudaFuncSetAttribute(selftest_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> hd(128 * 128), hr(128 * 128);
  float worst = 0.f;
  int failed = 0;
  int idx = 0;
  for (const SelfCase& c : cases) {
    ++idx;
    selftest_fill_kernel<<<64, 256>>>(A, 128 * c.K, 17u * idx);
    selftest_fill_kernel<<<64, 256>>>(B, c.N * c.K, 91u * idx + 5);
    cudaMemset(D, 0xff, 128 * 128 * 4);
    CUtensorMap tmA, tmB;
    memset(&tmA, 0, sizeof(tmA)); memset(&tmB, 0, sizeof(tmB));
    if (c.tma) {
      // 3-D maps with two trailing singleton dims keep one load helper for every case
      uint64_t da[3], db[3]; int64_t sa[3], sb[3]; uint32_t ba[3], bb[3];
      if (!c.a_mn) { da[0] = c.K; da[1] = 128; sa[1] = c.K; ba[0] = 64; ba[1] = 128; }
      else { da[0] = 128; da[1] = c.K; sa[1] = 128; ba[0] = 64; ba[1] = c.K; }
      if (!c.b_mn) { db[0] = c.K; db[1] = c.N; sb[1] = c.K; bb[0] = 64; bb[1] = c.N; }
      else { db[0] = c.N; db[1] = c.K; sb[1] = c.N; bb[0] = 64; bb[1] = c.K; }
      da[2] = db[2] = 1; sa[0] = sb[0] = 1; sa[2] = 128 * 128; sb[2] = 128 * 128; ba[2] = bb[2] = 1;
      int r1 = make_tmap_bf16(&tmA, A, 3, da, sa, ba, true);
      int r2 = make_tmap_bf16(&tmB, B, 3, db, sb, bb, true);
      if (r1 || r2) { failed = idx; break; }
    }
    selftest_umma_kernel<<<1, 128, smem>>>(A, B, D, c, tmA, tmB);
    selftest_ref_kernel<<<128, 128>>>(A, B, R, c);
    if (cudaDeviceSynchronize() != cudaSuccess) { g_last_cuda_error = (int)cudaGetLastError(); failed = idx; break; }
    cudaMemcpy(hd.data(), D, 128 * c.N * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(hr.data(), R, 128 * c.N * 4, cudaMemcpyDeviceToHost);
    float e = 0.f;
    for (int i = 0; i < 128 * c.N; ++i) {
      float d = fabsf(hd[i] - hr[i]);
      if (!(d <= 1e30f)) d = 1e30f;
      if (d > e) e = d;
    }
    if (e > worst) worst = e;
    if (e > 1e-3f) {
      fprintf(stderr, "[hymeta selftest] case %d (a_mn=%d b_mn=%d N=%d K=%d tma=%d) max err %g\n", idx,
              c.a_mn, c.b_mn, c.N, c.K, c.tma, e);
      if (!failed) failed = idx;
    }
  }
  cudaFree(A); cudaFree(B); cudaFree(D); cudaFree(R);
  if (max_err_out) *max_err_out = worst;
  count_launch((int)(sizeof(cases) / sizeof(cases[0])));
  return failed;
}
