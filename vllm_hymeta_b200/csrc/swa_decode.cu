// SWA + meta-prefix paged decode, split-K (a6/a8).  HBM-bound: the floor is one read of
// the visible K/V rows.  One launch (+ a tiny combine) replaces flash-attn fwd_kvcache + the meta
// pass + the 6-kernel TorchScript LSE merge of the reference (fa_kernel_metatoken.py:124-194).
//
// Grid (splits, Hkv * head_groups, B), 160 threads = 4 consumer warps + 1 TMA producer warp.
//   producer : one elected lane walks the split's 64-key tiles; per tile it resolves the page from
//              the block table ONCE (a tile never straddles a page: it is aligned to 64 logical
//              positions, or is cut into gcd(block_size, 64)-row boxes) and issues bulk-tensor
//              copies (cp.async.bulk.tensor, SWIZZLE_128B) of the K and V rows of ONE kv-head
//              straight out of the [num_blocks, block_size, Hkv, D] cache into a 3-stage smem ring
//              guarded by full/empty mbarriers.  No per-row address arithmetic anywhere.
//   consumers: warp w owns keys [16w, 16w+16) of every tile and serves all G = H/Hkv q-heads of the
//              GQA group from that single read (q-heads padded to the 16-row MMA tile): S = Q K^T
//              and O += P V on mma.sync m16n8k16 with an independent online softmax per warp; no
//              CTA-wide barrier inside the loop.  The four partial softmaxes are merged in smem and
//              the CTA emits a normalised partial (o, lse) that swa_decode_combine_kernel reduces.
// Split 0 is the shared meta prefix (always visible) when M > 0; the remaining splits partition the
// window [max(0, L-1-window_left), L) on 64-key boundaries of the LOGICAL position.
// Why mma.sync and not tcgen05: 7 q-heads x 2 x 128 MACs per 512 B of KV is ~90 TFLOP/s at the HBM
// roofline -- above the fp32 pipe, far below either tensor path; the kernel is judged on GB/s.
#include "umma.cuh"

#include <cstdlib>

namespace hymeta {

constexpr int TK = 64;                       // keys per tile
constexpr int NST = 3;                       // ring stages
constexpr int TILE_BYTES = TK * kD * 2;      // 16 KB per K (or V) tile
constexpr int SD_Q = 0;                                  // [16][128] bf16          4 KB
constexpr int SD_K = 4096;                               // [NST][2][64][64] bf16  48 KB
constexpr int SD_V = SD_K + NST * TILE_BYTES;            //                        48 KB
constexpr int SD_BAR = SD_V + NST * TILE_BYTES;          // full[NST], empty[NST]
constexpr int SWA_DEC_SMEM = SD_BAR + 64 + 1024;         // + slack for 1024 B alignment
constexpr int SD_THREADS = 160;

struct SwaDecodeParams {
  const __nv_bfloat16 *q;
  const int32_t *block_tables, *seqlens;
  float *part_o, *part_lse;
  __nv_bfloat16* out;
  int B, H, Hkv, M, block_size, window_left, n_splits, tps, n_hgroups, box_rows;
  float scale_log2;
  int64_t bt_stride, q_sb, q_sh, o_sb, o_sh;
};

// Q tile: [16 rows][128] bf16, 16-byte chunks XOR-swizzled per row (ldmatrix conflict free)
__device__ __forceinline__ uint32_t swz_q(int r, int c) { return (uint32_t)(r * 256 + ((c ^ (r & 7)) << 4)); }
// K/V tile as TMA lands it: two [64 rows][64 el] SWIZZLE_128B halves; c = 16-byte chunk 0..15 of the row
__device__ __forceinline__ uint32_t swz_t(int r, int c) {
  return (uint32_t)((c >> 3) * 8192 + r * 128 + (((c & 7) ^ (r & 7)) << 4));
}

__global__ void __launch_bounds__(SD_THREADS)
swa_decode_split_kernel(const SwaDecodeParams p, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmMK,
                        const __grid_constant__ CUtensorMap tmMV) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sbase = smem_u32(smem);
  const uint32_t sQ = sbase + SD_Q, sK = sbase + SD_K, sV = sbase + SD_V;
  const uint32_t bar_full = sbase + SD_BAR, bar_empty = bar_full + 8 * NST;

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

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < NST; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 4); }
    fence_mbar_init();
  }
  __syncthreads();   // barriers initialised; the producer starts streaming while Q is fetched

  if (warp == 4) {
    // ------------------------------------------------------------------ TMA producer (one lane)
    if (lane == 0) {
      const CUtensorMap* mk = is_meta ? &tmMK : &tmK;
      const CUtensorMap* mv = is_meta ? &tmMV : &tmV;
      tma_prefetch_desc(mk); tma_prefetch_desc(mv);
      const int rb = p.box_rows, nsub = TK / rb;
      const int32_t* bt = p.block_tables + (int64_t)b * p.bt_stride;
      const int last_pos = max(vhi - 1, 0);
      for (int it = 0; it < n_tiles; ++it) {
        const int st = it % NST;
        if (it >= NST) mbar_wait(bar_empty + 8 * st, ((it / NST) - 1) & 1);
        const uint32_t bar = bar_full + 8 * st;
        mbar_expect_tx(bar, 2 * TILE_BYTES);
        const uint32_t dk = sK + st * TILE_BYTES, dv = sV + st * TILE_BYTES;
        const int pos0 = (t0 + it) * TK;
        for (int sub = 0; sub < nsub; ++sub) {
          int row, blk;
          if (is_meta) { row = pos0 + sub * rb; blk = 0; }   // rows >= M: TMA zero-fills
          else {
            // boxes wholly past the sequence are redirected to the page of the last valid position
            // (their rows are masked; V rows >= L are zeroed by the consumers)
            // likewise boxes wholly before the window start go to the box holding `lo`
            const int pos = min(max(pos0 + sub * rb, vlo / rb * rb), last_pos);
            blk = bt[pos / p.block_size];
            row = (pos % p.block_size) / rb * rb;
          }
          const uint32_t o = sub * rb * 128;
          tma_load_4d(dk + o, mk, bar, 0, hkv, row, blk);
          tma_load_4d(dk + 8192 + o, mk, bar, 64, hkv, row, blk);
          tma_load_4d(dv + o, mv, bar, 0, hkv, row, blk);
          tma_load_4d(dv + 8192 + o, mv, bar, 64, hkv, row, blk);
        }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumers (warps 0..3)
  const int g = lane >> 2, t4 = lane & 3;   // accumulator row / column-pair owner
  const int lm = lane >> 3, lr = lane & 7;  // ldmatrix: matrix index / row inside it
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  float o_acc[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) o_acc[i][j] = 0.f;
  {  // Q: 16 rows x 16 chunks, rows >= ng zero-filled
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int idx = tid + i * 128, r = idx >> 4, c = idx & 15;
      const bool ok = r < ng;
      const __nv_bfloat16* src = p.q + b * p.q_sb + (int64_t)(h0 + (ok ? r : 0)) * p.q_sh + c * 8;
      cp_async16(sQ + swz_q(r, c), src, ok);
    }
    cp_async_commit();
    cp_async_wait<0>();
    asm volatile("bar.sync 1, 128;" ::: "memory");
  }
  uint32_t qa[8][4];
#pragma unroll
  for (int ks = 0; ks < 8; ++ks)
    ldmatrix_x4(qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3], sQ + swz_q((lm & 1) * 8 + lr, 2 * ks + (lm >> 1)));

  for (int it = 0; it < n_tiles; ++it) {
    const int st = it % NST;
    mbar_wait(bar_full + 8 * st, (it / NST) & 1);
    const uint32_t kst = sK + st * TILE_BYTES, vst = sV + st * TILE_BYTES;
    const int pos0 = (t0 + it) * TK + 16 * warp;

    if (pos0 < vlo || pos0 + 16 > vhi) {
      // rows outside [vlo, vhi) may hold uninitialised / recycled cache bytes (NaN patterns):
      // P = 0 there, but 0 * NaN would poison O, so clear this warp's masked V rows
      const int r = 16 * warp + (lane >> 1);
      const int rp = (t0 + it) * TK + r;
      if (rp < vlo || rp >= vhi) {
        uint8_t* row = smem + SD_V + st * TILE_BYTES + (lane & 1) * 8192 + r * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(row + c * 16) = make_uint4(0, 0, 0, 0);
      }
      __syncwarp();
    }

    // S[16 heads x 16 keys] = Q K^T for this warp's 16 keys
    float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      uint32_t b0, b1, b2, b3;
      ldmatrix_x4(b0, b1, b2, b3, kst + swz_t(16 * warp + ((lm >> 1) << 3) + lr, 2 * ks + (lm & 1)));
      mma_bf16_16816(s[0], qa[ks], b0, b1);
      mma_bf16_16816(s[1], qa[ks], b2, b3);
    }
    // mask + scale (log2 domain)
    if (pos0 < vlo || pos0 + 16 > vhi) {
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int pos = pos0 + nt * 8 + 2 * t4 + (e & 1);
          s[nt][e] = (pos >= vlo && pos < vhi) ? s[nt][e] * p.scale_log2 : -INFINITY;
        }
    } else {
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) s[nt][e] *= p.scale_log2;
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
    if (__any_sync(0xffffffffu, alpha[0] != 1.f || alpha[1] != 1.f)) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        o_acc[i][0] *= alpha[0]; o_acc[i][1] *= alpha[0];
        o_acc[i][2] *= alpha[1]; o_acc[i][3] *= alpha[1];
      }
    }
    uint32_t pa[4];
    pa[0] = pack_bf16(ps[0][0], ps[0][1]); pa[1] = pack_bf16(ps[0][2], ps[0][3]);
    pa[2] = pack_bf16(ps[1][0], ps[1][1]); pa[3] = pack_bf16(ps[1][2], ps[1][3]);
    // O[16 heads x 128] += P[16 x 16 keys] V[16 keys x 128]
#pragma unroll
    for (int dn = 0; dn < 8; ++dn) {
      uint32_t v0, v1, v2, v3;
      ldmatrix_x4_trans(v0, v1, v2, v3, vst + swz_t(16 * warp + ((lm & 1) << 3) + lr, 2 * dn + (lm >> 1)));
      mma_bf16_16816(o_acc[2 * dn], pa, v0, v1);
      mma_bf16_16816(o_acc[2 * dn + 1], pa, v2, v3);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_empty + 8 * st);   // this warp is done with stage st
  }

  // ---- merge the four warps (smem reuse: every tile has landed and been consumed) ----------------
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    l_run[rr] += __shfl_xor_sync(0xffffffffu, l_run[rr], 1);
    l_run[rr] += __shfl_xor_sync(0xffffffffu, l_run[rr], 2);
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");   // all consumer warps are out of the ring
  constexpr int OS = kD + 4;  // padded row stride (floats)
  float* sO = reinterpret_cast<float*>(smem + SD_K);  // [4][16][OS]
  float* sM = sO + 4 * 16 * OS;                        // [4][16]
  float* sL = sM + 64;                                 // [4][16]
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    *reinterpret_cast<float2*>(&sO[(warp * 16 + g) * OS + i * 8 + 2 * t4]) = make_float2(o_acc[i][0], o_acc[i][1]);
    *reinterpret_cast<float2*>(&sO[(warp * 16 + g + 8) * OS + i * 8 + 2 * t4]) = make_float2(o_acc[i][2], o_acc[i][3]);
  }
  if (t4 == 0) {
    sM[warp * 16 + g] = m_run[0]; sM[warp * 16 + g + 8] = m_run[1];
    sL[warp * 16 + g] = l_run[0]; sL[warp * 16 + g + 8] = l_run[1];
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");
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

static int env_int(const char* name, int dflt) {
  const char* s = getenv(name);
  return s && *s ? atoi(s) : dflt;
}

static SwaDecodePlan plan_swa_decode(int B, int H, int Hkv, int max_blocks, int block_size,
                                     int window_left, int M) {
  SwaDecodePlan pl;
  const int G = H / Hkv;
  pl.n_hgroups = (G + 15) / 16;
  long max_keys = (long)max_blocks * block_size;
  if (window_left >= 0 && window_left + 1 < max_keys) max_keys = window_left + 1;
  if (max_keys < 1) max_keys = 1;
  const int max_tiles = (int)((max_keys + TK - 1) / TK) + 1;  // +1: window start not tile aligned
  const long units = (long)B * Hkv * pl.n_hgroups;
  // CTAs of ~8 tiles (256 KB of K/V) keep the tail of the last wave short; fewer tiles per CTA
  // only when the batch alone cannot fill 148 SMs x 2 CTAs a few times over
  static const int tps_env = env_int("HYMETA_SWA_DEC_TPS", 0);   // tuning knob (0 = heuristic)
  long tps = tps_env > 0 ? tps_env : 8;
  if (tps_env <= 0) {
    const long want = 148L * 2 * 4;   // >= 4 waves of CTAs
    while (tps > 1 && units * ((max_tiles + tps - 1) / tps) < want) tps >>= 1;
  }
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
    size_t workspace_bytes, int B, int H, int Hkv, int D, int M, int block_size, int num_blocks,
    int max_blocks, int64_t bt_stride, int window_left, float scale, int64_t q_sb, int64_t q_sh,
    int64_t o_sb, int64_t o_sh, int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk,
    int64_t vc_stok, int64_t vc_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
    void* stream) {
  if (B == 0) return HYMETA_OK;
  if (!q || !k_cache || !v_cache || !block_tables || !cache_seqlens || !out || !workspace || B < 0 ||
      H <= 0 || Hkv <= 0 || M < 0 || block_size <= 0 || max_blocks < 0 || num_blocks <= 0)
    return HYMETA_ERR_INVALID_ARG;
  if (M > 0 && (!meta_k || !meta_v)) return HYMETA_ERR_INVALID_ARG;
  if (D != kD || H % Hkv) return HYMETA_ERR_UNSUPPORTED;
  if (block_size % 16) return HYMETA_ERR_UNSUPPORTED;   // reference: block_size % 16 == 0 (meta_attention.py:88-91)
  if (B > 65535) return HYMETA_ERR_UNSUPPORTED;
  const int64_t strides[] = {q_sb, q_sh, kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh,
                             M ? mk_st : 0, M ? mk_sh : 0, M ? mv_st : 0, M ? mv_sh : 0};
  for (int64_t s : strides)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;
  const SwaDecodePlan pl = plan_swa_decode(B, H, Hkv, max_blocks, block_size, window_left, M);
  if (workspace_bytes < pl.bytes) return HYMETA_ERR_WORKSPACE;

  SwaDecodeParams p;
  p.q = (const __nv_bfloat16*)q; p.block_tables = block_tables; p.seqlens = cache_seqlens;
  p.part_o = (float*)workspace;
  p.part_lse = p.part_o + (size_t)B * H * pl.n_splits * kD;
  p.out = (__nv_bfloat16*)out;
  p.B = B; p.H = H; p.Hkv = Hkv; p.M = M; p.block_size = block_size; p.window_left = window_left;
  p.n_splits = pl.n_splits; p.tps = pl.tps; p.n_hgroups = pl.n_hgroups;
  p.box_rows = block_size % 64 == 0 ? 64 : (block_size % 32 == 0 ? 32 : 16);
  p.scale_log2 = scale * 1.4426950408889634f;
  p.bt_stride = bt_stride; p.q_sb = q_sb; p.q_sh = q_sh; p.o_sb = o_sb; p.o_sh = o_sh;

  // tensor maps: [D, Hkv, block_size, num_blocks] over the paged caches, [D, Hkv, M, 1] over meta
  CUtensorMap tmK, tmV, tmMK, tmMV;
  auto mk4 = [&](CUtensorMap* tm, const void* base, int rows, int blocks, int64_t s_h, int64_t s_tok,
                 int64_t s_blk, uint32_t box_rows) {
    const uint64_t dims[4] = {(uint64_t)kD, (uint64_t)Hkv, (uint64_t)rows, (uint64_t)blocks};
    const int64_t str[4] = {1, s_h, s_tok, s_blk};
    const uint32_t box[4] = {64, 1, box_rows, 1};
    return make_tmap_bf16(tm, base, 4, dims, str, box, true);
  };
  int rc;
  if ((rc = mk4(&tmK, k_cache, block_size, num_blocks, kc_sh, kc_stok, kc_sblk, p.box_rows))) return rc;
  if ((rc = mk4(&tmV, v_cache, block_size, num_blocks, vc_sh, vc_stok, vc_sblk, p.box_rows))) return rc;
  if (M > 0) {
    // meta rows use the same box height as the cache so the producer loop is uniform
    if ((rc = mk4(&tmMK, meta_k, M, 1, mk_sh, mk_st, (int64_t)M * mk_st, p.box_rows))) return rc;
    if ((rc = mk4(&tmMV, meta_v, M, 1, mv_sh, mv_st, (int64_t)M * mv_st, p.box_rows))) return rc;
  } else {
    tmMK = tmK; tmMV = tmV;
  }

  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(swa_decode_split_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         SWA_DEC_SMEM);
    attr_set = true;
  }
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(pl.n_splits, Hkv * pl.n_hgroups, B);
  swa_decode_split_kernel<<<grid, SD_THREADS, SWA_DEC_SMEM, st>>>(p, tmK, tmV, tmMK, tmMV);
  swa_decode_combine_kernel<<<dim3(H, B), 128, 0, st>>>(p);
  count_launch(2);
  return check_launch();
}
