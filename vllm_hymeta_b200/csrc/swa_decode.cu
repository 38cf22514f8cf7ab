// SWA + meta-prefix paged decode, split-K (a6/a8).  HBM-bound: the floor is one read of
// the visible K/V rows.  One launch replaces flash-attn fwd_kvcache + the meta pass + the 6-kernel
// TorchScript LSE merge of the reference (fa_kernel_metatoken.py:124-194).
//
// Grid (splits, Hkv * head_groups, B), 128 threads.  A CTA streams 64-key tiles of ONE kv-head
// through a 3-stage cp.async ring (swizzled smem) and serves all G = H/Hkv q-heads of the group
// from that single read (q-heads padded to the 16-row MMA tile).  The four warps each own 16 keys
// of a tile and run an independent online softmax; they are merged in smem at the end and the
// CTA emits a normalised partial (o, lse) that swa_decode_combine_kernel reduces over splits.
// Split 0 is the shared meta prefix (always visible) when M > 0; the remaining splits partition
// the window [max(0, L-1-window_left), L) on 64-key boundaries of the LOGICAL position, so a tile
// never straddles more pages than block_size allows and every block-table lookup is per row.
// The math is CUDA-core-proof: 7 q-heads x 2 x 128 MACs per 512 B of KV needs ~90 TFLOP/s at the
// HBM roofline, above the fp32 pipe, hence mma.sync here (tensor rate is irrelevant: see DESIGN.md).
#include "common.cuh"

namespace hymeta {

constexpr int TK = 64;                       // keys per tile
constexpr int NST = 3;                       // cp.async stages
constexpr int TILE_BYTES = TK * kD * 2;      // 16 KB
constexpr int SWA_DEC_SMEM = 16 * kD * 2 + 2 * NST * TILE_BYTES;  // Q + K/V rings = 100 KB

struct SwaDecodeParams {
  const __nv_bfloat16 *q, *k_cache, *v_cache, *meta_k, *meta_v;
  const int32_t *block_tables, *seqlens;
  float *part_o, *part_lse;
  __nv_bfloat16* out;
  int B, H, Hkv, M, block_size, window_left, n_splits, tps, n_hgroups;
  float scale_log2;
  int64_t bt_stride, q_sb, q_sh, o_sb, o_sh, kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh,
      mk_st, mk_sh, mv_st, mv_sh;
};

// byte offset of 16-byte chunk `c` of row `r` in a [rows][128] bf16 tile, XOR-swizzled so that the
// 8 rows of an ldmatrix 8x8 block land in 8 different bank groups
__device__ __forceinline__ uint32_t swz(int r, int c) { return (uint32_t)(r * 256 + ((c ^ (r & 7)) << 4)); }

__global__ void __launch_bounds__(128) swa_decode_split_kernel(const SwaDecodeParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t sQ = smem_u32(smem);
  const uint32_t sK = sQ + 16 * kD * 2;
  const uint32_t sV = sK + NST * TILE_BYTES;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int split = blockIdx.x;
  const int hkv = blockIdx.y / p.n_hgroups, hg = blockIdx.y % p.n_hgroups;
  const int b = blockIdx.z;
  const int G = p.H / p.Hkv;
  const int ng = min(16, G - hg * 16);
  const int h0 = hkv * G + hg * 16;

  const int L = p.seqlens[b];
  const int lo = p.window_left < 0 ? 0 : max(0, L - 1 - p.window_left);
  const bool is_meta = p.M > 0 && split == 0;
  int t0, t1, vlo, vhi;  // tile range and valid position range
  if (is_meta) {
    t0 = 0; t1 = (p.M + TK - 1) / TK; vlo = 0; vhi = p.M;
  } else {
    const int rs = split - (p.M > 0 ? 1 : 0);
    t0 = lo / TK + rs * p.tps;
    t1 = min(t0 + p.tps, (L + TK - 1) / TK);
    vlo = lo; vhi = L;
  }
  const int n_tiles = max(0, t1 - t0);
  const int64_t part_base = ((int64_t)b * p.H + h0) * p.n_splits + split;  // + row * n_splits
  if (n_tiles == 0) {
    if (tid < ng) p.part_lse[part_base + (int64_t)tid * p.n_splits] = -INFINITY;
    return;
  }

  // ---- loaders ---------------------------------------------------------------------------------
  {  // Q: 16 rows x 16 chunks, rows >= ng zero-filled
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * 128, r = idx >> 4, c = idx & 15;
      const bool ok = r < ng;
      const __nv_bfloat16* src = p.q + b * p.q_sb + (int64_t)(h0 + (ok ? r : 0)) * p.q_sh + c * 8;
      cp_async16(sQ + swz(r, c), src, ok);
    }
  }
  auto load_tile = [&](int stage, int t) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int idx = tid + i * 128, r = idx >> 4, c = idx & 15;
      const int pos = t * TK + r;
      const bool ok = pos >= vlo && pos < vhi;
      const __nv_bfloat16 *ks, *vs;
      if (is_meta) {
        const int pp = ok ? pos : 0;
        ks = p.meta_k + pp * p.mk_st + hkv * p.mk_sh + c * 8;
        vs = p.meta_v + pp * p.mv_st + hkv * p.mv_sh + c * 8;
      } else {
        int64_t blk = 0, off = 0;
        if (ok) {
          blk = p.block_tables[b * p.bt_stride + pos / p.block_size];
          off = pos % p.block_size;
        }
        ks = p.k_cache + blk * p.kc_sblk + off * p.kc_stok + hkv * p.kc_sh + c * 8;
        vs = p.v_cache + blk * p.vc_sblk + off * p.vc_stok + hkv * p.vc_sh + c * 8;
      }
      cp_async16(sK + stage * TILE_BYTES + swz(r, c), ks, ok);
      cp_async16(sV + stage * TILE_BYTES + swz(r, c), vs, ok);
    }
  };

#pragma unroll
  for (int s = 0; s < NST - 1; ++s) {
    if (s < n_tiles) load_tile(s, t0 + s);
    cp_async_commit();
  }

  // ---- per-warp online softmax state -----------------------------------------------------------
  const int g = lane >> 2, t4 = lane & 3;   // accumulator row / column-pair owner
  const int lm = lane >> 3, lr = lane & 7;  // ldmatrix: matrix index / row inside it
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  float o_acc[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) o_acc[i][j] = 0.f;
  uint32_t qa[8][4];

  for (int it = 0; it < n_tiles; ++it) {
    if (it + NST - 1 < n_tiles) load_tile((it + NST - 1) % NST, t0 + it + NST - 1);
    cp_async_commit();
    cp_async_wait<NST - 1>();
    __syncthreads();
    if (it == 0) {
#pragma unroll
      for (int ks = 0; ks < 8; ++ks)
        ldmatrix_x4(qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3],
                    sQ + swz((lm & 1) * 8 + lr, 2 * ks + (lm >> 1)));
    }
    const uint32_t kst = sK + (it % NST) * TILE_BYTES, vst = sV + (it % NST) * TILE_BYTES;

    // S[16 heads x 16 keys] = Q K^T for this warp's 16 keys
    float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      uint32_t b0, b1, b2, b3;
      ldmatrix_x4(b0, b1, b2, b3, kst + swz(16 * warp + ((lm >> 1) << 3) + lr, 2 * ks + (lm & 1)));
      mma_bf16_16816(s[0], qa[ks], b0, b1);
      mma_bf16_16816(s[1], qa[ks], b2, b3);
    }
    // mask + scale (log2 domain)
    const int pos0 = (t0 + it) * TK + 16 * warp;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int pos = pos0 + nt * 8 + 2 * t4 + (e & 1);
        s[nt][e] = (pos >= vlo && pos < vhi) ? s[nt][e] * p.scale_log2 : -INFINITY;
      }
    float alpha[2], msafe[2];
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      float mx = fmaxf(fmaxf(s[0][2 * rr], s[0][2 * rr + 1]), fmaxf(s[1][2 * rr], s[1][2 * rr + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float mn = fmaxf(m_run[rr], mx);
      alpha[rr] = (mn == -INFINITY) ? 1.f : exp2f(m_run[rr] - mn);
      msafe[rr] = (mn == -INFINITY) ? 0.f : mn;
      m_run[rr] = mn;
    }
    float ps[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) ps[nt][e] = exp2f(s[nt][e] - msafe[e >> 1]);
#pragma unroll
    for (int rr = 0; rr < 2; ++rr)
      l_run[rr] = l_run[rr] * alpha[rr] + ps[0][2 * rr] + ps[0][2 * rr + 1] + ps[1][2 * rr] + ps[1][2 * rr + 1];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      o_acc[i][0] *= alpha[0]; o_acc[i][1] *= alpha[0];
      o_acc[i][2] *= alpha[1]; o_acc[i][3] *= alpha[1];
    }
    uint32_t pa[4];
    pa[0] = pack_bf16(ps[0][0], ps[0][1]); pa[1] = pack_bf16(ps[0][2], ps[0][3]);
    pa[2] = pack_bf16(ps[1][0], ps[1][1]); pa[3] = pack_bf16(ps[1][2], ps[1][3]);
    // O[16 heads x 128] += P[16 x 16 keys] V[16 keys x 128]
#pragma unroll
    for (int dn = 0; dn < 8; ++dn) {
      uint32_t v0, v1, v2, v3;
      ldmatrix_x4_trans(v0, v1, v2, v3, vst + swz(16 * warp + ((lm & 1) << 3) + lr, 2 * dn + (lm >> 1)));
      mma_bf16_16816(o_acc[2 * dn], pa, v0, v1);
      mma_bf16_16816(o_acc[2 * dn + 1], pa, v2, v3);
    }
    __syncthreads();  // stage may be overwritten by the next iteration's prefetch
  }
  cp_async_wait<0>();

  // ---- merge the four warps (smem reuse: the rings are idle now) ---------------------------------
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    l_run[rr] += __shfl_xor_sync(0xffffffffu, l_run[rr], 1);
    l_run[rr] += __shfl_xor_sync(0xffffffffu, l_run[rr], 2);
  }
  constexpr int OS = kD + 4;  // padded row stride (floats)
  float* sO = reinterpret_cast<float*>(smem + 16 * kD * 2);  // [4][16][OS]
  float* sM = sO + 4 * 16 * OS;                               // [4][16]
  float* sL = sM + 64;                                        // [4][16]
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    *reinterpret_cast<float2*>(&sO[(warp * 16 + g) * OS + i * 8 + 2 * t4]) = make_float2(o_acc[i][0], o_acc[i][1]);
    *reinterpret_cast<float2*>(&sO[(warp * 16 + g + 8) * OS + i * 8 + 2 * t4]) = make_float2(o_acc[i][2], o_acc[i][3]);
  }
  if (t4 == 0) {
    sM[warp * 16 + g] = m_run[0]; sM[warp * 16 + g + 8] = m_run[1];
    sL[warp * 16 + g] = l_run[0]; sL[warp * 16 + g + 8] = l_run[1];
  }
  __syncthreads();
  for (int r = 0; r < ng; ++r) {
    const float mm = fmaxf(fmaxf(sM[r], sM[16 + r]), fmaxf(sM[32 + r], sM[48 + r]));
    const int64_t pi = part_base + (int64_t)r * p.n_splits;
    if (mm == -INFINITY) {
      if (tid == 0) p.part_lse[pi] = -INFINITY;
      continue;
    }
    float lt = 0.f, acc = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float wt = exp2f(sM[w * 16 + r] - mm);
      lt += wt * sL[w * 16 + r];
      acc += wt * sO[(w * 16 + r) * OS + tid];
    }
    p.part_o[pi * kD + tid] = acc / lt;
    if (tid == 0) p.part_lse[pi] = mm + log2f(lt);
  }
}

// out[b,h,:] = sum_s w_s part_o[b,h,s,:],  w_s = 2^(lse_s - max) / sum  (grid (H, B), 128 threads)
__global__ void __launch_bounds__(128) swa_decode_combine_kernel(const SwaDecodeParams p) {
  const int h = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
  const int64_t base = ((int64_t)b * p.H + h) * p.n_splits;
  float mm = -INFINITY;
  for (int s = 0; s < p.n_splits; ++s) mm = fmaxf(mm, p.part_lse[base + s]);
  float acc = 0.f, wsum = 0.f;
  if (mm != -INFINITY) {
    for (int s = 0; s < p.n_splits; ++s) {
      const float ls = p.part_lse[base + s];
      if (ls == -INFINITY) continue;
      const float w = exp2f(ls - mm);
      wsum += w;
      acc += w * p.part_o[(base + s) * kD + d];
    }
    acc /= wsum;
  }
  p.out[b * p.o_sb + h * p.o_sh + d] = __float2bfloat16_rn(acc);
}

struct SwaDecodePlan { int n_splits, tps, n_hgroups; size_t bytes; };

static SwaDecodePlan plan_swa_decode(int B, int H, int Hkv, int max_blocks, int block_size,
                                     int window_left, int M) {
  SwaDecodePlan pl;
  const int G = H / Hkv;
  pl.n_hgroups = (G + 15) / 16;
  long max_keys = (long)max_blocks * block_size;
  if (window_left >= 0 && window_left + 1 < max_keys) max_keys = window_left + 1;
  if (max_keys < 1) max_keys = 1;
  const int max_tiles = (int)((max_keys + TK - 1) / TK) + 1;  // +1: window start not tile aligned
  long units = (long)B * Hkv * pl.n_hgroups;
  long tps = units * max_tiles / 1184;  // aim for >= ~8 CTAs per SM before giving up splits
  if (tps < 1) tps = 1;
  if (tps > max_tiles) tps = max_tiles;
  pl.tps = (int)tps;
  pl.n_splits = (max_tiles + pl.tps - 1) / pl.tps + (M > 0 ? 1 : 0);
  pl.bytes = (size_t)B * H * pl.n_splits * (kD + 1) * sizeof(float);
  return pl;
}

}  // namespace hymeta

using namespace hymeta;

extern "C" size_t hymeta_swa_decode_workspace_bytes(int B, int H, int Hkv, int D, int max_blocks,
                                                    int block_size, int window_left, int M) {
  if (B <= 0 || H <= 0 || Hkv <= 0 || H % Hkv || D != kD || block_size <= 0) return 0;
  return plan_swa_decode(B, H, Hkv, max_blocks, block_size, window_left, M).bytes;
}

extern "C" int hymeta_swa_meta_decode_paged(
    const void* q, const void* k_cache, const void* v_cache, const int32_t* block_tables,
    const int32_t* cache_seqlens, const void* meta_k, const void* meta_v, void* out, void* workspace,
    size_t workspace_bytes, int B, int H, int Hkv, int D, int M, int block_size, int max_blocks,
    int64_t bt_stride, int window_left, float scale, int64_t q_sb, int64_t q_sh, int64_t o_sb,
    int64_t o_sh, int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk, int64_t vc_stok,
    int64_t vc_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, void* stream) {
  if (B == 0) return HYMETA_OK;
  if (!q || !k_cache || !v_cache || !block_tables || !cache_seqlens || !out || !workspace || B < 0 ||
      H <= 0 || Hkv <= 0 || M < 0 || block_size <= 0 || max_blocks < 0)
    return HYMETA_ERR_INVALID_ARG;
  if (M > 0 && (!meta_k || !meta_v)) return HYMETA_ERR_INVALID_ARG;
  if (D != kD || H % Hkv) return HYMETA_ERR_UNSUPPORTED;
  if (B > 65535) return HYMETA_ERR_UNSUPPORTED;
  const int64_t strides[] = {q_sb, q_sh, kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh,
                             M ? mk_st : 0, M ? mk_sh : 0, M ? mv_st : 0, M ? mv_sh : 0};
  for (int64_t s : strides)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;
  const SwaDecodePlan pl = plan_swa_decode(B, H, Hkv, max_blocks, block_size, window_left, M);
  if (workspace_bytes < pl.bytes) return HYMETA_ERR_WORKSPACE;

  SwaDecodeParams p;
  p.q = (const __nv_bfloat16*)q; p.k_cache = (const __nv_bfloat16*)k_cache;
  p.v_cache = (const __nv_bfloat16*)v_cache; p.meta_k = (const __nv_bfloat16*)meta_k;
  p.meta_v = (const __nv_bfloat16*)meta_v; p.block_tables = block_tables; p.seqlens = cache_seqlens;
  p.part_o = (float*)workspace;
  p.part_lse = p.part_o + (size_t)B * H * pl.n_splits * kD;
  p.out = (__nv_bfloat16*)out;
  p.B = B; p.H = H; p.Hkv = Hkv; p.M = M; p.block_size = block_size; p.window_left = window_left;
  p.n_splits = pl.n_splits; p.tps = pl.tps; p.n_hgroups = pl.n_hgroups;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.bt_stride = bt_stride; p.q_sb = q_sb; p.q_sh = q_sh; p.o_sb = o_sb; p.o_sh = o_sh;
  p.kc_sblk = kc_sblk; p.kc_stok = kc_stok; p.kc_sh = kc_sh; p.vc_sblk = vc_sblk; p.vc_stok = vc_stok;
  p.vc_sh = vc_sh; p.mk_st = mk_st; p.mk_sh = mk_sh; p.mv_st = mv_st; p.mv_sh = mv_sh;

  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(swa_decode_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         SWA_DEC_SMEM);
    attr_set = true;
  }
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(pl.n_splits, Hkv * pl.n_hgroups, B);
  swa_decode_split_kernel<<<grid, 128, SWA_DEC_SMEM, st>>>(p);
  swa_decode_combine_kernel<<<dim3(H, B), 128, 0, st>>>(p);
  count_launch(2);
  return check_launch();
}
