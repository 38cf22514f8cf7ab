// GLA recurrent decode step (a1/a2) -- HBM-bound read-modify-write of the
// per-(request, state-head) 128x128 bf16 state.
//
//   S <- bf16(k (x) v) + exp(g)[:,None] * S        (fp32 math, bf16 store)
//   o  = (q * scale)^T S                            (all q-heads sharing the state)
//
// Persistent grid (<= 2 CTAs of 256 threads per SM), item = one (request, state-head) state; item
// i+2 streams into a 3-stage shared-memory ring with cp.async while item i is computed (see the
// kernel comment).  Warp w owns d_k rows [16w, 16w+16), a lane owns 8 consecutive d_v columns, so
// every state byte is touched by exactly one thread and stored back from registers with 128-bit
// stores.  The G = H/Hs q-heads of a GQA group are all served from the single state read.
// Instruction diet (the kernel is issue-bound long before it is FLOP-bound: 7 heads x 128 x 128 FMAs
// per 32 KB of state): the bf16 product k*v is ONE packed HMUL2.BF16 per column pair (bit-identical
// to the reference's bf16 `b_k[:, None] * b_v[None, :]`, my_fused_recurrent.py:65), the decay and the
// per-head read-out run on packed FFMA2 (fma.rn.f32x2), q is fetched as two 128-bit shared loads
// per row for all heads.  The reduction over d_k is log2(rows per warp instr) shuffle stages plus
// one pass through shared memory across the 4 warps.
#include "common.cuh"

#include <cstdlib>

// libdevice's accurate single-precision functions (what expf / log1pf resolve to without fast-math)
extern "C" __device__ float __nv_expf(float);
extern "C" __device__ float __nv_log1pf(float);

namespace hymeta {

struct GlaDecodeParams {
  const __nv_bfloat16 *q, *k, *v;
  const void* g;
  __nv_bfloat16* o;
  __nv_bfloat16* state;
  const int32_t* slot_idx;
  int B, H, Hk, Hv, Hs;
  uint32_t kmagic, vmagic;   // (hs * magic) >> 16 == hs / (Hs/Hk) resp. hs / (Hs/Hv), checked on the host
  int64_t q_sb, q_sh, k_sb, k_sh, v_sb, v_sh, g_sb, g_sh, o_sb, o_sh, st_ss, st_sh;
  float scale;
  const __nv_bfloat16* gnorm_w;   // per-head RMSNorm weight [128] of the 7B g_norm epilogue, or null
  float gnorm_eps;
  // > 0: q, k and g are the RAW projection outputs of the 7B layer and the feature map + gate
  // (model_for_7B/gla_attention.py:251-262) run here: q = relu(q), k = relu(k),
  // g = bf16(bf16(logsigmoid(gk)) / gate_norm) -- bit-identical to hymeta_gla_gate_7b's outputs: this
  // file is compiled with --use_fast_math, so the gate calls libdevice's accurate expf / log1pf by
  // name (flush-to-zero only matters where exp(-|x|) is denormal, and there log1p of it vanishes
  // against x either way); gate_norm must be a power of two so that / gate_norm == * gate_inv exactly
  float gate_norm, gate_inv;
};

typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi) {
  u64 r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(u64 v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) {
  u64 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint32_t hmul2_bf16(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

constexpr int GD_NS = 3;                         // ring stages (one item each)
constexpr int GD_STATE_BYTES = kD * kD * 2;       // 32 KB state
constexpr int GD_RAW_Q = GD_STATE_BYTES;          // raw q   [8 heads][128] bf16   2 KB
constexpr int GD_RAW_K = GD_RAW_Q + 2048;         // raw k   [128] bf16
constexpr int GD_RAW_V = GD_RAW_K + 256;          // raw v   [128] bf16
constexpr int GD_RAW_G = GD_RAW_V + 256;          // raw g   [128] fp32 (or bf16 in the first half)
constexpr int GD_STAGE = GD_RAW_G + 512;          // 35840 B per stage
constexpr int GD_THREADS = 256;
constexpr int GD_SQ = GD_NS * GD_STAGE;           // s_q  [128][8] f32  (scaled, fp32)
constexpr int GD_SKG = GD_SQ + kD * 8 * 4;        // s_kg [128] {k bf16x2, exp(g)}
constexpr int GD_SLOT = GD_SKG + kD * 8;          // slot ring [8] int
constexpr int GD_SMEM = GD_SLOT + 32;

// Persistent, software-pipelined: CTA c handles items c, c+grid, ... (item = one (request,
// state-head) 128x128 state with its small operands).  Item i+2 streams into a 3-stage
// shared-memory ring with cp.async while item i is being computed, so no global-load latency is
// ever exposed inside the loop: per item the CTA waits on one cp.async group, converts the small
// operands warp-locally (each warp only needs q/k/g of its own 16 d_k rows), runs the FFMA2 main
// loop out of shared memory and stores the new state straight from registers.
// GT >= G = H/Hs: q-heads served per state (accumulators per lane: GT x 8 columns).
template <int GT, bool GF32, bool GN>
__global__ void __launch_bounds__(GD_THREADS, 2) gla_decode_kernel(const GlaDecodeParams p, const int n_items) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ float s_norm[32];                      // g_norm epilogue: [head][warp of the half-CTA] sums of squares
  __shared__ float s_gw[GN ? kD : 1];               // g_norm weight, staged once per CTA (a per-item global load
                                                    // sits exposed on the epilogue's critical path)
  float (*s_q)[8] = reinterpret_cast<float (*)[8]>(smem + GD_SQ);
  uint2* s_kg = reinterpret_cast<uint2*>(smem + GD_SKG);
  const uint32_t ring = smem_u32(smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int c = lane & 15, rg = lane >> 4;          // 16-byte column chunk / row inside a warp instr
  const int G = p.H / p.Hs;
  const int gs = gridDim.x;
  pdl_wait();
  pdl_launch_dependents();
  if (GN && tid < kD) s_gw[tid] = __bfloat162float(p.gnorm_w[tid]);   // visible after the first __syncthreads
  const int gs_b = gs / p.Hs, gs_h = gs % p.Hs;     // item stride decomposed: no per-item division

  // Item bookkeeping without registers held across the loop: (b, hs) of items k, k+2, k+4 advance
  // incrementally; the slot index of item k+4 is fetched with a 4-byte cp.async into a small ring
  // in the same commit group as item k+2's operands, so it is visible (wait_group + barrier) two
  // iterations before issue() needs it and no thread ever stalls on the slot_idx load.
  int* s_slot = reinterpret_cast<int*>(smem + GD_SLOT);
  struct Item { int b, hs; };
  auto advance = [&](Item x) -> Item {              // item + gs
    x.b += gs_b; x.hs += gs_h;
    if (x.hs >= p.Hs) { x.hs -= p.Hs; ++x.b; }
    return x;
  };
  // small-operand copy role of this thread: one 16-byte chunk (or none)
  //   tid <  128: q head tid/16, rows 8*(tid%16)..   128..143: k   144..159: v   160..191: g
  auto issue = [&](const Item& x, int slot, int stage) {
    if (slot >= 0) {
      const uint32_t sb = ring + stage * GD_STAGE;
      {  // this thread's 8 chunks of the state
        const __nv_bfloat16* src = p.state + (int64_t)slot * p.st_ss + (int64_t)x.hs * p.st_sh +
                                   (warp * 16 + rg) * kD + c * 8;
        const uint32_t dst = sb + (warp * 16 + rg) * 256 + c * 16;
#pragma unroll
        for (int i = 0; i < 8; ++i) cp_async16(dst + i * 512, src + i * 2 * kD, true);
      }
      if (tid < 128) {
        const int hh = tid >> 4, ch = tid & 15;
        const bool ok = hh < G;
        cp_async16(sb + GD_RAW_Q + tid * 16, p.q + x.b * p.q_sb + (int64_t)(x.hs * G + (ok ? hh : 0)) * p.q_sh + ch * 8, ok);
      } else if (tid < 144) {
        cp_async16(sb + GD_RAW_K + (tid - 128) * 16,
                   p.k + x.b * p.k_sb + (int64_t)((x.hs * p.kmagic) >> 16) * p.k_sh + (tid - 128) * 8, true);
      } else if (tid < 160) {
        cp_async16(sb + GD_RAW_V + (tid - 144) * 16,
                   p.v + x.b * p.v_sb + (int64_t)((x.hs * p.vmagic) >> 16) * p.v_sh + (tid - 144) * 8, true);
      } else if (tid < (GF32 ? 192 : 176)) {
        const int64_t off = x.b * p.g_sb + (int64_t)((x.hs * p.kmagic) >> 16) * p.g_sh;
        if (GF32) cp_async16(sb + GD_RAW_G + (tid - 160) * 16, reinterpret_cast<const float*>(p.g) + off + (tid - 160) * 4, true);
        else cp_async16(sb + GD_RAW_G + (tid - 160) * 16, reinterpret_cast<const __nv_bfloat16*>(p.g) + off + (tid - 160) * 8, true);
      }
    }
  };
  auto fetch_slot = [&](const Item& x, int j) {     // one thread: slot of item j -> ring entry j & 7
    if (x.b < p.B) cp_async4(smem_u32(s_slot + (j & 7)), p.slot_idx + x.b);
    else s_slot[j & 7] = -1;
  };

  Item cur_i;
  cur_i.b = blockIdx.x / p.Hs; cur_i.hs = blockIdx.x % p.Hs;
  Item i2 = advance(advance(cur_i)), i4 = advance(advance(i2));
  if (tid < 4) {                                     // slots of items 0..3, once, synchronously
    Item x = cur_i;
    for (int j = 0; j < tid; ++j) x = advance(x);
    s_slot[tid] = x.b < p.B ? p.slot_idx[x.b] : -1;
  }
  __syncthreads();
  issue(cur_i, s_slot[0], 0);
  cp_async_commit();
  issue(advance(cur_i), s_slot[1], 1);
  cp_async_commit();

  for (int k = 0; cur_i.b < p.B; ++k) {
    const int stage = k % GD_NS;
    cp_async_wait<1>();                              // own chunks of item k have landed ...
    __syncthreads();                                 // ... and everyone else's; item k-1 is fully consumed
    issue(i2, s_slot[(k + 2) & 7], (k + 2) % GD_NS);  // refill the stage item k-1 used
    if (tid == GD_THREADS - 1) fetch_slot(i4, k + 4);
    cp_async_commit();
    const int cur_slot = s_slot[k & 7];
    const int h0 = cur_i.hs * G;
    __nv_bfloat16* op = p.o + cur_i.b * p.o_sb + (int64_t)(h0 + (tid >> 7)) * p.o_sh + (tid & 127);
    if (cur_slot < 0) {   // PAD slot (CUDA-graph padding): skip, emit zeros (CTA-uniform branch)
#pragma unroll 1
      for (int hh = tid >> 7; hh < G; hh += 2, op += 2 * p.o_sh) *op = __float2bfloat16(0.f);
      cur_i = advance(cur_i); i2 = advance(i2); i4 = advance(i4);
      continue;
    }
    uint8_t* sbase = smem + stage * GD_STAGE;
    {  // warp-local conversion of the small operands of this warp's 16 rows
      const int row = warp * 16 + (lane & 15), hf = lane >> 4;
      const __nv_bfloat16* rq = reinterpret_cast<const __nv_bfloat16*>(sbase + GD_RAW_Q) + row;
      const bool gate = p.gate_norm > 0.f;
      const float qlo = gate ? 0.f : -INFINITY;     // relu as a max against 0 (identity against -inf)
      float4 t;
      t.x = fmaxf(__bfloat162float(rq[(hf * 4 + 0) * kD]), qlo) * p.scale;
      t.y = fmaxf(__bfloat162float(rq[(hf * 4 + 1) * kD]), qlo) * p.scale;
      t.z = fmaxf(__bfloat162float(rq[(hf * 4 + 2) * kD]), qlo) * p.scale;
      t.w = fmaxf(__bfloat162float(rq[(hf * 4 + 3) * kD]), qlo) * p.scale;
      *reinterpret_cast<float4*>(&s_q[row][hf * 4]) = t;
      if (hf == 0) {
        uint32_t kb = (uint32_t)reinterpret_cast<const uint16_t*>(sbase + GD_RAW_K)[row];
        if (gate && (kb & 0x8000u)) kb = 0u;       // relu(k) on the bf16 bits
        s_kg[row].x = kb | (kb << 16);
      } else {
        float gv;
        if (GF32) gv = reinterpret_cast<const float*>(sbase + GD_RAW_G)[row];
        else gv = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(sbase + GD_RAW_G)[row]);
        if (!GF32 && gate) {
          // logsigmoid(x) = min(x, 0) - log1p(exp(-|x|)) (ATen), rounded to bf16 before the division
          // and again after it -- exactly what hymeta_gla_gate_7b stores (gla_gate.cu)
          const float ls = fminf(gv, 0.f) - __nv_log1pf(__nv_expf(-fabsf(gv)));
          // gate_norm is a power of two (checked on the host): the division is an exact multiply, and
          // the kernel stays free of the IEEE-division slow-path call (its mere presence cost the
          // main loop 23 % at B = 1000)
          gv = round_bf16(round_bf16(ls) * p.gate_inv);
        }
        s_kg[row].y = __float_as_uint(__expf(gv));
      }
    }
    // Main loop: a lane owns 4 consecutive d_v columns (8 bytes per row) of ALL 16 rows of its warp,
    // so a warp instruction moves one whole 256-byte state row and the d_k reduction inside a warp
    // is the lane's own accumulation.  (With 8 columns per lane and the rows split over the two
    // half-warps, the epilogue needed 64 shuffles per thread -- at one warp-shuffle per clock per SM
    // that was a third of the per-item time at B = 64 -- and 8 x GT packed accumulators spilled.)
    const uint2 vraw = *reinterpret_cast<const uint2*>(sbase + GD_RAW_V + lane * 8);
    const uint32_t vv[2] = {vraw.x, vraw.y};
    __nv_bfloat16* sp = p.state + (int64_t)cur_slot * p.st_ss + (int64_t)cur_i.hs * p.st_sh +
                        (warp * 16) * kD + lane * 4;
    const uint32_t st_base = ring + stage * GD_STAGE + (warp * 16) * 256 + lane * 8;
    __syncwarp();

    u64 acc[GT][2];
#pragma unroll
    for (int hh = 0; hh < GT; ++hh) { acc[hh][0] = 0ull; acc[hh][1] = 0ull; }

    const uint2* kgp = s_kg + warp * 16;
    const float* qrow = &s_q[warp * 16][0];
#pragma unroll
    for (int i = 0; i < 16; ++i) {                   // row r = warp*16 + i
      uint2 w2;
      asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(w2.x), "=r"(w2.y) : "r"(st_base + i * 256));
      const uint2 kg = kgp[i];
      const float eg = __uint_as_float(kg.y);
      const uint32_t w[2] = {w2.x, w2.y};
      u64 sn[2];   // new state, fp32 pairs, un-rounded (the output reduces this, my_fused_recurrent.py:75)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const uint32_t kv = hmul2_bf16(kg.x, vv[j]);            // bf16(k * v), both columns
        sn[j] = ffma2(pack2(eg, eg), pack2(bf16_lo(w[j]), bf16_hi(w[j])),
                      pack2(bf16_lo(kv), bf16_hi(kv)));
      }
      {
        uint2 st;
        float a, bb;
        unpack2(sn[0], a, bb); st.x = pack_bf16(a, bb);
        unpack2(sn[1], a, bb); st.y = pack_bf16(a, bb);
        asm volatile("st.global.L1::no_allocate.v2.u32 [%0], {%1,%2};"
                     :: "l"(sp + i * kD), "r"(st.x), "r"(st.y) : "memory");
      }
      float qr[8];
      {
        const float4 t0 = *reinterpret_cast<const float4*>(qrow + i * 8);
        qr[0] = t0.x; qr[1] = t0.y; qr[2] = t0.z; qr[3] = t0.w;
        if (GT > 4) {
          const float4 t1 = *reinterpret_cast<const float4*>(qrow + i * 8 + 4);
          qr[4] = t1.x; qr[5] = t1.y; qr[6] = t1.z; qr[7] = t1.w;
        }
      }
#pragma unroll
      for (int hh = 0; hh < GT; ++hh) {
        const u64 q2 = pack2(qr[hh], qr[hh]);
        acc[hh][0] = ffma2(q2, sn[0], acc[hh][0]);
        acc[hh][1] = ffma2(q2, sn[1], acc[hh][1]);
      }
    }

    // reduce over d_k across the 8 warps through smem
    // (the scratch aliases this item's state: every thread is done reading it after the barrier)
    __syncthreads();
    float* s_red = reinterpret_cast<float*>(sbase);   // [8][GT][128]
#pragma unroll
    for (int hh = 0; hh < GT; ++hh) {
      float4 a;
      unpack2(acc[hh][0], a.x, a.y);
      unpack2(acc[hh][1], a.z, a.w);
      *reinterpret_cast<float4*>(&s_red[(warp * GT + hh) * kD + lane * 4]) = a;
    }
    __syncthreads();
    if (GN) {
      // 7B g_norm (model_for_7B/gla_attention.py:177,233): vLLM RMSNorm over the head's 128 outputs
      // of the bf16-rounded o: y = bf16(bf16(o * rstd) * w).  The 128 threads of a half-CTA hold one
      // head at a time (up to 4 heads each): all sums of squares are reduced first (warp shuffle,
      // 4 warp partials per head in smem), ONE 128-thread named barrier, then scale and store.
      const float* rp = s_red + (tid >> 7) * kD + (tid & 127);
      float xs[4];
#pragma unroll
      for (int n = 0; n < 4; ++n) {
        const int hh = (tid >> 7) + 2 * n;
        xs[n] = 0.f;
        if (hh < G) {
          float x = 0.f;
#pragma unroll
          for (int w = 0; w < 8; ++w) x += rp[(2 * n * kD) + w * GT * kD];
          const float xb = round_bf16(x);
          xs[n] = xb;
          float ss = xb * xb;
#pragma unroll
          for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
          if (lane == 0) s_norm[hh * 4 + ((tid >> 5) & 3)] = ss;
        }
      }
      if (tid >> 7) asm volatile("bar.sync 2, 128;" ::: "memory");
      else asm volatile("bar.sync 1, 128;" ::: "memory");
      const float gw = s_gw[tid & 127];
#pragma unroll
      for (int n = 0; n < 4; ++n) {
        const int hh = (tid >> 7) + 2 * n;
        if (hh < G) {
          const float* sn = s_norm + hh * 4;
          const float tot = (sn[0] + sn[1]) + (sn[2] + sn[3]);
          op[(int64_t)2 * n * p.o_sh] =
              __float2bfloat16_rn(round_bf16(xs[n] * rsqrtf(tot * (1.f / kD) + p.gnorm_eps)) * gw);
        }
      }
    } else {
      const float* rp = s_red + (tid >> 7) * kD + (tid & 127);
#pragma unroll 1
      for (int hh = tid >> 7; hh < G; hh += 2, op += 2 * p.o_sh, rp += 2 * kD) {
        float x = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) x += rp[w * GT * kD];
        *op = __float2bfloat16_rn(x);
      }
    }
    // no barrier here: the next iteration's top barrier orders these reads before the refill
    cur_i = advance(cur_i); i2 = advance(i2); i4 = advance(i4);
  }
  cp_async_wait<0>();
}

template <int GT>
static void launch_gla_decode(const GlaDecodeParams& p, bool g_fp32, cudaStream_t st) {
  const long n_items = (long)p.B * p.Hs;
  const int grid = (int)(n_items < 2 * 148 ? n_items : 2 * 148);
  static DynSmemAttr attr;   // one per template instantiation of this launcher
  if (attr.need()) {
    DynSmemAttr::set(gla_decode_kernel<GT, true, false>, GD_SMEM);
    DynSmemAttr::set(gla_decode_kernel<GT, false, false>, GD_SMEM);
    DynSmemAttr::set(gla_decode_kernel<GT, true, true>, GD_SMEM);
    DynSmemAttr::set(gla_decode_kernel<GT, false, true>, GD_SMEM);
    attr.mark();
  }
  const dim3 gr(grid), bl(GD_THREADS);
  const int n = (int)n_items;
  if (p.gnorm_w) {
    if (g_fp32) launch_pdl(gla_decode_kernel<GT, true, true>, gr, bl, GD_SMEM, st, p, n);
    else launch_pdl(gla_decode_kernel<GT, false, true>, gr, bl, GD_SMEM, st, p, n);
  } else {
    if (g_fp32) launch_pdl(gla_decode_kernel<GT, true, false>, gr, bl, GD_SMEM, st, p, n);
    else launch_pdl(gla_decode_kernel<GT, false, false>, gr, bl, GD_SMEM, st, p, n);
  }
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_gla_decode_gate7b(const void* q, const void* k, const void* v, const void* g, void* o,
                                 void* state, const int32_t* slot_idx, int B, int H, int Hk, int Hv,
                                 int Hs, int D, int64_t q_sb, int64_t q_sh, int64_t k_sb, int64_t k_sh,
                                 int64_t v_sb, int64_t v_sh, int64_t g_sb, int64_t g_sh,
                                 int64_t o_sb, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                                 int g_is_fp32, float scale, const void* gnorm_weight, float gnorm_eps,
                                 float gate_normalizer, void* stream) {
  if (B == 0) return HYMETA_OK;
  if (gate_normalizer < 0.f || (gate_normalizer > 0.f && g_is_fp32)) return HYMETA_ERR_INVALID_ARG;
  if (gate_normalizer > 0.f) {   // power of two only (the model's 16): see the kernel
    int e = 0;
    if (frexpf(gate_normalizer, &e) != 0.5f) return HYMETA_ERR_UNSUPPORTED;
  }
  if (!q || !k || !v || !g || !o || !state || !slot_idx || B < 0 || H <= 0 || Hk <= 0 || Hv <= 0 || Hs <= 0)
    return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  if (H % Hs != 0 || Hs % Hk != 0 || Hs % Hv != 0) return HYMETA_ERR_INVALID_ARG;
  if (H / Hs > 8 || Hs > 4096) return HYMETA_ERR_UNSUPPORTED;   // as the prefill kernel: at most 8 q-heads per state
  // every operand is moved with 16-byte cp.async: bases and strides must keep that alignment
  const int gdiv = g_is_fp32 ? 4 : 8;
  if ((st_ss % 8) || (st_sh % 8) || (v_sb % 8) || (v_sh % 8) || (q_sb % 8) || (q_sh % 8) || (k_sb % 8) ||
      (k_sh % 8) || (g_sb % gdiv) || (g_sh % gdiv))
    return HYMETA_ERR_INVALID_ARG;
  if (((uintptr_t)q | (uintptr_t)k | (uintptr_t)v | (uintptr_t)g | (uintptr_t)state) & 15)
    return HYMETA_ERR_INVALID_ARG;
  if (B > 65535) return HYMETA_ERR_UNSUPPORTED;
  GlaDecodeParams p{(const __nv_bfloat16*)q, (const __nv_bfloat16*)k, (const __nv_bfloat16*)v, g,
                    (__nv_bfloat16*)o, (__nv_bfloat16*)state, slot_idx, B, H, Hk, Hv, Hs, 0u, 0u,
                    q_sb, q_sh, k_sb, k_sh, v_sb, v_sh, g_sb, g_sh, o_sb, o_sh, st_ss, st_sh, scale,
                    (const __nv_bfloat16*)gnorm_weight, gnorm_eps, gate_normalizer,
                    gate_normalizer > 0.f ? 1.f / gate_normalizer : 0.f};
  auto magic = [&](int div) -> uint32_t {   // exact for hs < Hs <= 4096 (verified below)
    const uint32_t m = (65536u + div - 1) / div;
    for (int x = 0; x < Hs; ++x)
      if ((int)((x * m) >> 16) != x / div) return 0u;
    return m;
  };
  p.kmagic = magic(Hs / Hk); p.vmagic = magic(Hs / Hv);
  if (!p.kmagic || !p.vmagic) return HYMETA_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  const int G = H / Hs;
  if ((long)B * Hs > 0x7fffffffL) return HYMETA_ERR_UNSUPPORTED;
  const bool f = g_is_fp32 != 0;
  if (G == 1) launch_gla_decode<1>(p, f, st);
  else if (G <= 4) launch_gla_decode<4>(p, f, st);
  else if (G == 7) launch_gla_decode<7>(p, f, st);   // spill-free since a lane holds 2 (not 4) packed accumulators per head
  else launch_gla_decode<8>(p, f, st);
  count_launch();
  return check_launch();
}

extern "C" int hymeta_gla_decode_gnorm(const void* q, const void* k, const void* v, const void* g, void* o,
                                 void* state, const int32_t* slot_idx, int B, int H, int Hk, int Hv,
                                 int Hs, int D, int64_t q_sb, int64_t q_sh, int64_t k_sb, int64_t k_sh,
                                 int64_t v_sb, int64_t v_sh, int64_t g_sb, int64_t g_sh,
                                 int64_t o_sb, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                                 int g_is_fp32, float scale, const void* gnorm_weight, float gnorm_eps,
                                 void* stream) {
  return hymeta_gla_decode_gate7b(q, k, v, g, o, state, slot_idx, B, H, Hk, Hv, Hs, D, q_sb, q_sh, k_sb,
                                  k_sh, v_sb, v_sh, g_sb, g_sh, o_sb, o_sh, st_ss, st_sh, g_is_fp32,
                                  scale, gnorm_weight, gnorm_eps, 0.f, stream);
}

extern "C" int hymeta_gla_decode(const void* q, const void* k, const void* v, const void* g, void* o,
                                 void* state, const int32_t* slot_idx, int B, int H, int Hk, int Hv,
                                 int Hs, int D, int64_t q_sb, int64_t q_sh, int64_t k_sb, int64_t k_sh,
                                 int64_t v_sb, int64_t v_sh, int64_t g_sb, int64_t g_sh,
                                 int64_t o_sb, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                                 int g_is_fp32, float scale, void* stream) {
  return hymeta_gla_decode_gnorm(q, k, v, g, o, state, slot_idx, B, H, Hk, Hv, Hs, D, q_sb, q_sh, k_sb,
                                 k_sh, v_sb, v_sh, g_sb, g_sh, o_sb, o_sh, st_ss, st_sh, g_is_fp32,
                                 scale, nullptr, 0.f, stream);
}
