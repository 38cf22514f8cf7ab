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
  // fused paged-KV write of the step's new token (position L-1): null = the pages already hold it
  const __nv_bfloat16 *new_k, *new_v;
  __nv_bfloat16 *k_cache, *v_cache;
  int64_t nk_sb, nk_sh, nv_sb, nv_sh, kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh;
  // slot of each request's new token (reshape_and_cache_flash convention: < 0 = padded row, no
  // store); null = the page of position L-1 taken from the block table
  const int64_t* new_slots;
  // fused combine: one ticket per (request, kv-head group); null = separate combine kernel
  int* tickets;
};

__device__ __forceinline__ float ldcg_f32(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

__device__ __forceinline__ float4 ldcg_f32x4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}

// Split-K combine for the `ng` q-heads h0.. of one (request, kv-head group), run by 128 threads of
// whichever split CTA finishes last (ticket), or by the combine kernel.  Partials written by other
// CTAs are read through L2 (ld.cg).  Latency-lean: all lse values are fetched in one round trip,
// warp w then owns heads w, w+4, ... and keeps every partial-output load of a head in flight at
// once (lane = 4 columns).  Splits are summed in index order: the result does not depend on which
// CTA runs it.  s_buf: ng * n_splits floats.
__device__ __forceinline__ void swa_combine_heads(const SwaDecodeParams& p, int b, int h0, int ng, int tid,
                                                  float* s_buf) {
  const int lane = tid & 31, warp = tid >> 5, ns = p.n_splits;
  const int64_t base0 = ((int64_t)b * p.H + h0) * ns;
  const int total = ng * ns;
  for (int i = tid; i < total; i += 128) s_buf[i] = ldcg_f32(p.part_lse + base0 + i);
  asm volatile("bar.sync 1, 128;" ::: "memory");
  for (int r = warp; r < ng; r += 4) {
    float* w = s_buf + r * ns;
    float mx = -INFINITY;
    for (int s = lane; s < ns; s += 32) mx = fmaxf(mx, w[s]);
#pragma unroll
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int s = lane; s < ns; s += 32) {
      const float e = mx == -INFINITY ? 0.f : exp2f(w[s] - mx);
      w[s] = e;
      sum += e;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    __syncwarp();
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* po = p.part_o + (base0 + (int64_t)r * ns) * kD + lane * 4;
#pragma unroll 8
    for (int s = 0; s < ns; ++s) {
      const float e = w[s];
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);     // empty splits never wrote their partial
      if (e != 0.f) o = ldcg_f32x4(po + (int64_t)s * kD);
      acc.x = fmaf(e, o.x, acc.x); acc.y = fmaf(e, o.y, acc.y);
      acc.z = fmaf(e, o.z, acc.z); acc.w = fmaf(e, o.w, acc.w);
    }
    const float inv = sum > 0.f ? 1.f / sum : 0.f;
    uint2 pk;
    pk.x = pack_bf16(acc.x * inv, acc.y * inv);
    pk.y = pack_bf16(acc.z * inv, acc.w * inv);
    *reinterpret_cast<uint2*>(p.out + b * p.o_sb + (int64_t)(h0 + r) * p.o_sh + lane * 4) = pk;
  }
}

// Ticket: the partials of this CTA are written; returns true in the CTA that arrives last for its
// (request, kv-head group) -- every other split's partials are then visible to it.  128 threads.
__device__ __forceinline__ bool swa_take_ticket(const SwaDecodeParams& p, int b, int tid, int* s_flag) {
  __threadfence();                                         // publish this thread's partial stores
  asm volatile("bar.sync 1, 128;" ::: "memory");
  if (tid == 0) {
    int* t = p.tickets + (int64_t)b * gridDim.y + blockIdx.y;
    const int old = atomicAdd(t, 1);
    const int last = old == (int)gridDim.x - 1;
    if (last) *t = 0;                                      // restore for the next launch
    *s_flag = last;
    __threadfence();
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");
  return *s_flag != 0;
}

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
  pdl_wait();                 // seqlens / pages / q come from the kernels ahead in the stream
  pdl_launch_dependents();
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
    if (p.n_splits == 1) {     // single split: this CTA owns the output rows (no visible key: zeros)
      if (tid < 128)
        for (int r = 0; r < ng; ++r) p.out[b * p.o_sb + (int64_t)(h0 + r) * p.o_sh + tid] = __float2bfloat16(0.f);
      return;
    }
    if (tid < ng) p.part_lse[part_base + (int64_t)tid * p.n_splits] = -INFINITY;
    if (p.tickets && tid < 128) {   // (n_splits > 1 here)
      float* s_buf = reinterpret_cast<float*>(smem + SD_K);
      if (swa_take_ticket(p, b, tid, reinterpret_cast<int*>(smem + SD_Q)))
        swa_combine_heads(p, b, h0, ng, tid, s_buf);
    }
    return;
  }
  // the step's new token (fused paged-KV write): it lives in tile (L-1)/TK of exactly one real split
  const int new_it = (p.new_k && !is_meta) ? (L - 1) / TK - t0 : -1;
  const bool use_new = new_it >= 0 && new_it < n_tiles;

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

  // new-token row: warp (r / 16) patches smem row r = (L-1) % TK of the last tile with new_k/new_v
  // (lanes 0-15: K chunks, 16-31: V chunks) and, for head group 0, writes the row to its page
  const int new_r = (L - 1) % TK;
  const bool new_mine = use_new && warp == (new_r >> 4);
  uint4 new_x = make_uint4(0, 0, 0, 0);
  if (new_mine) {
    const __nv_bfloat16* src = lane < 16 ? p.new_k + b * p.nk_sb + (int64_t)hkv * p.nk_sh
                                         : p.new_v + b * p.nv_sb + (int64_t)hkv * p.nv_sh;
    new_x = ldg_nc_v4(src + (lane & 15) * 8);
    int64_t blk;
    int off;
    bool store = hg == 0;
    if (p.new_slots) {
      const int64_t slot = p.new_slots[b];
      store = store && slot >= 0;
      blk = slot / p.block_size; off = (int)(slot % p.block_size);
    } else {
      const int pos = L - 1;
      blk = p.block_tables[(int64_t)b * p.bt_stride + pos / p.block_size];
      off = pos % p.block_size;
    }
    if (store) {
      __nv_bfloat16* dst = lane < 16 ? p.k_cache + blk * p.kc_sblk + off * p.kc_stok + (int64_t)hkv * p.kc_sh
                                     : p.v_cache + blk * p.vc_sblk + off * p.vc_stok + (int64_t)hkv * p.vc_sh;
      stg_v4(dst + (lane & 15) * 8, new_x);
    }
  }

  for (int it = 0; it < n_tiles; ++it) {
    const int st = it % NST;
    mbar_wait(bar_full + 8 * st, (it / NST) & 1);
    const uint32_t kst = sK + st * TILE_BYTES, vst = sV + st * TILE_BYTES;
    const int pos0 = (t0 + it) * TK + 16 * warp;
    if (new_mine && it == new_it) {
      *reinterpret_cast<uint4*>(smem + (lane < 16 ? SD_K : SD_V) + st * TILE_BYTES + swz_t(new_r, lane & 15)) = new_x;
      __syncwarp();
    }

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
    __nv_bfloat16* orow = p.out + b * p.o_sb + (int64_t)(h0 + r) * p.o_sh;
    if (mm == -INFINITY) {
      if (p.n_splits == 1) orow[tid] = __float2bfloat16(0.f);
      else if (tid == 0) p.part_lse[pi] = -INFINITY;
      continue;
    }
    float lt = 0.f, acc = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float wt = exp2f(sM[w * 16 + r] - mm);
      lt += wt * sL[w * 16 + r];
      acc += wt * sO[(w * 16 + r) * OS + tid];
    }
    if (p.n_splits == 1) { orow[tid] = __float2bfloat16_rn(acc / lt); continue; }   // final: no combine pass
    p.part_o[pi * kD + tid] = acc / lt;
    if (tid == 0) p.part_lse[pi] = mm + log2f(lt);
  }
  if (p.tickets && p.n_splits > 1) {
    asm volatile("bar.sync 1, 128;" ::: "memory");        // sM / sL / sO fully consumed
    float* s_buf = reinterpret_cast<float*>(smem + SD_K);   // ring is idle: 96 KB >= 16 heads x 1024 splits
    if (swa_take_ticket(p, b, tid, reinterpret_cast<int*>(smem + SD_Q)))
      swa_combine_heads(p, b, h0, ng, tid, s_buf);
  }
}

// Separate combine launch (tickets == null): the same routine, grid (Hkv * head_groups, B), 128 threads
__global__ void __launch_bounds__(128) swa_decode_combine_kernel(const SwaDecodeParams p) {
  extern __shared__ float s_comb[];
  pdl_wait();
  pdl_launch_dependents();
  const int hkv = blockIdx.x / p.n_hgroups, hg = blockIdx.x % p.n_hgroups, b = blockIdx.y;
  const int G = p.H / p.Hkv;
  swa_combine_heads(p, b, hkv * G + hg * 16, min(16, G - hg * 16), threadIdx.x, s_comb);
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
  // Tiles per CTA, from B200 sweeps (tools/ab_swa_decode.py, B = 1 .. 256 inside a CUDA graph): every
  // wave of CTAs pays ~2 tile-times of ramp / drain that the co-resident CTA does not hide (the
  // CTAs of a wave start and end in lock-step), so the best plan is ONE wave of long CTAs: as many
  // splits as it takes to fill the 2 x 148 CTA slots, never CTAs shorter than 8 tiles (256 KB of
  // K/V), and a single split per (request, kv-head) once the batch fills the machine by itself.
  // Config 2 (256 units): 87.7 us per layer-step with one 66-tile CTA per unit against 104.7 us
  // with nine 8-tile splits.
  static const int tps_env = env_int("HYMETA_SWA_DEC_TPS", 0);   // tuning knob (0 = rule above)
  long tps = tps_env;
  if (tps_env <= 0) {
    long ns = 296 / (units > 0 ? units : 1);
    if (ns > max_tiles / 8) ns = max_tiles / 8;
    if (ns < 1) ns = 1;
    tps = (max_tiles + ns - 1) / ns;
  }
  if (tps > max_tiles) tps = max_tiles;
  while ((max_tiles + tps - 1) / tps > 1024) tps *= 2;   // the combine keeps 16 x n_splits weights in smem
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

extern "C" int hymeta_swa_meta_decode_paged_fused(
    const void* q, void* k_cache, void* v_cache, const int32_t* block_tables,
    const int32_t* cache_seqlens, const void* meta_k, const void* meta_v, void* out, void* workspace,
    size_t workspace_bytes, int B, int H, int Hkv, int D, int M, int block_size, int num_blocks,
    int max_blocks, int64_t bt_stride, int window_left, float scale, int64_t q_sb, int64_t q_sh,
    int64_t o_sb, int64_t o_sh, int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk,
    int64_t vc_stok, int64_t vc_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
    const void* new_k, const void* new_v, int64_t nk_sb, int64_t nk_sh, int64_t nv_sb, int64_t nv_sh,
    const int64_t* new_slot_mapping, int32_t* tickets, void* stream) {
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
  if ((new_k == nullptr) != (new_v == nullptr)) return HYMETA_ERR_INVALID_ARG;
  if (new_k && ((nk_sb % 8) || (nk_sh % 8) || (nv_sb % 8) || (nv_sh % 8) ||
                (((uintptr_t)new_k | (uintptr_t)new_v) & 15)))
    return HYMETA_ERR_INVALID_ARG;
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
  p.new_k = (const __nv_bfloat16*)new_k; p.new_v = (const __nv_bfloat16*)new_v;
  p.k_cache = (__nv_bfloat16*)k_cache; p.v_cache = (__nv_bfloat16*)v_cache;
  p.nk_sb = nk_sb; p.nk_sh = nk_sh; p.nv_sb = nv_sb; p.nv_sh = nv_sh;
  p.kc_sblk = kc_sblk; p.kc_stok = kc_stok; p.kc_sh = kc_sh;
  p.vc_sblk = vc_sblk; p.vc_stok = vc_stok; p.vc_sh = vc_sh;
  p.tickets = tickets;
  p.new_slots = new_k ? new_slot_mapping : nullptr;

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

  static DynSmemAttr attr;
  if (attr.need()) {
    DynSmemAttr::set(swa_decode_split_kernel, SWA_DEC_SMEM);
    DynSmemAttr::set(swa_decode_combine_kernel, 16 * 1025 * (int)sizeof(float));
    attr.mark();
  }
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(pl.n_splits, Hkv * pl.n_hgroups, B);
  launch_pdl(swa_decode_split_kernel, grid, dim3(SD_THREADS), SWA_DEC_SMEM, st, p, tmK, tmV, tmMK, tmMV);
  if (!tickets && pl.n_splits > 1)
    launch_pdl(swa_decode_combine_kernel, dim3(Hkv * pl.n_hgroups, B), dim3(128),
               (size_t)16 * pl.n_splits * sizeof(float), st, p);
  count_launch(tickets || pl.n_splits == 1 ? 1 : 2);
  return check_launch();
}

extern "C" size_t hymeta_swa_decode_ticket_bytes(int B, int H, int Hkv) {
  if (B <= 0 || H <= 0 || Hkv <= 0 || H % Hkv) return 0;
  return (size_t)B * Hkv * ((H / Hkv + 15) / 16) * sizeof(int32_t);
}

extern "C" int hymeta_swa_meta_decode_paged(
    const void* q, const void* k_cache, const void* v_cache, const int32_t* block_tables,
    const int32_t* cache_seqlens, const void* meta_k, const void* meta_v, void* out, void* workspace,
    size_t workspace_bytes, int B, int H, int Hkv, int D, int M, int block_size, int num_blocks,
    int max_blocks, int64_t bt_stride, int window_left, float scale, int64_t q_sb, int64_t q_sh,
    int64_t o_sb, int64_t o_sh, int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk,
    int64_t vc_stok, int64_t vc_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
    void* stream) {
  return hymeta_swa_meta_decode_paged_fused(
      q, const_cast<void*>(k_cache), const_cast<void*>(v_cache), block_tables, cache_seqlens, meta_k,
      meta_v, out, workspace, workspace_bytes, B, H, Hkv, D, M, block_size, num_blocks, max_blocks,
      bt_stride, window_left, scale, q_sb, q_sh, o_sb, o_sh, kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok,
      vc_sh, mk_st, mk_sh, mv_st, mv_sh, nullptr, nullptr, 0, 0, 0, 0, nullptr, nullptr, stream);
}

// ---- no-allocation variant -----------------------------------------------------------------------
// For callers that cannot provide (or allocate) the split-K scratch: it lives in this file-scope
// device array -- HYMETA_SWA_DECODE_STATIC_WS_BYTES of fp32 partials followed by
// HYMETA_SWA_DECODE_STATIC_TICKET_BYTES of tickets -- which every GPU that loads the library holds
// until exit.  In host code the array's name is only a handle: its device address is looked up once
// per device (with that device current) through cudaGetSymbolAddress and passed on like a
// caller-owned workspace.
static __device__ __align__(256) unsigned char
    g_swa_decode_static_scratch[HYMETA_SWA_DECODE_STATIC_WS_BYTES + HYMETA_SWA_DECODE_STATIC_TICKET_BYTES];

static int swa_decode_static_scratch(unsigned char** out) {
  static std::atomic<unsigned char*> cache[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { g_last_cuda_error = (int)e; return HYMETA_ERR_CUDA; }
  if (dev < 0 || dev >= 64) return HYMETA_ERR_UNSUPPORTED;
  unsigned char* ptr = cache[dev].load(std::memory_order_acquire);
  if (!ptr) {
    e = cudaGetSymbolAddress((void**)&ptr, g_swa_decode_static_scratch);
    if (e != cudaSuccess || !ptr) { g_last_cuda_error = (int)e; return HYMETA_ERR_CUDA; }
    cache[dev].store(ptr, std::memory_order_release);
  }
  *out = ptr;
  return HYMETA_OK;
}

extern "C" int hymeta_swa_meta_decode_paged_static(
    const void* q, const void* k_cache, const void* v_cache, const int32_t* block_tables,
    const int32_t* cache_seqlens, const void* meta_k, const void* meta_v, void* out,
    int B, int H, int Hkv, int D, int M, int block_size, int num_blocks,
    int max_blocks, int64_t bt_stride, int window_left, float scale, int64_t q_sb, int64_t q_sh,
    int64_t o_sb, int64_t o_sh, int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk,
    int64_t vc_stok, int64_t vc_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
    void* stream) {
  if (B == 0) return HYMETA_OK;
  if (B < 0 || H <= 0 || Hkv <= 0 || block_size <= 0 || max_blocks < 0 || M < 0) return HYMETA_ERR_INVALID_ARG;
  if (D != kD || H % Hkv) return HYMETA_ERR_UNSUPPORTED;
  // the same plan as the caller-workspace entry points; sizes beyond the static array are rejected
  const size_t ws_bytes = plan_swa_decode(B, H, Hkv, max_blocks, block_size, window_left, M).bytes;
  const size_t ticket_bytes = hymeta_swa_decode_ticket_bytes(B, H, Hkv);
  if (ws_bytes > HYMETA_SWA_DECODE_STATIC_WS_BYTES || ticket_bytes > HYMETA_SWA_DECODE_STATIC_TICKET_BYTES)
    return HYMETA_ERR_WORKSPACE;
  unsigned char* scratch = nullptr;
  int rc = swa_decode_static_scratch(&scratch);
  if (rc) return rc;
  int32_t* tickets = reinterpret_cast<int32_t*>(scratch + HYMETA_SWA_DECODE_STATIC_WS_BYTES);
  // The array keeps the previous call's contents.  The partials are fully rewritten before they are
  // read; the tickets must start at zero (the kernel returns them to zero, but a call that was
  // aborted must not poison the next one), so they are cleared in stream order.
  cudaError_t e = cudaMemsetAsync(tickets, 0, ticket_bytes, (cudaStream_t)stream);
  if (e != cudaSuccess) { g_last_cuda_error = (int)e; return HYMETA_ERR_CUDA; }
  return hymeta_swa_meta_decode_paged_fused(
      q, const_cast<void*>(k_cache), const_cast<void*>(v_cache), block_tables, cache_seqlens, meta_k,
      meta_v, out, scratch, HYMETA_SWA_DECODE_STATIC_WS_BYTES, B, H, Hkv, D, M, block_size, num_blocks,
      max_blocks, bt_stride, window_left, scale, q_sb, q_sh, o_sb, o_sh, kc_sblk, kc_stok, kc_sh, vc_sblk,
      vc_stok, vc_sh, mk_st, mk_sh, mv_st, mv_sh, nullptr, nullptr, 0, 0, 0, 0, nullptr, tickets, stream);
}
