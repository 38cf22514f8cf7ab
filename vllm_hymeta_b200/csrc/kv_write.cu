// Paged-KV write (a9) and slot-state maintenance (a5).  Pure HBM copies, bit-exact.
#include "common.cuh"

namespace hymeta {

// One thread moves one 16-byte chunk of K and the matching chunk of V.
__global__ void __launch_bounds__(256)
kv_write_kernel(const __nv_bfloat16* __restrict__ key, const __nv_bfloat16* __restrict__ value,
                __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ v_cache,
                const int64_t* __restrict__ slot_mapping, int T, int Hkv, int block_size,
                int64_t key_st, int64_t key_sh, int64_t val_st, int64_t val_sh, int64_t kc_sblk,
                int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk, int64_t vc_stok, int64_t vc_sh) {
  constexpr int CPR = kD / 8;  // 16-byte chunks per head row
  pdl_wait();
  pdl_launch_dependents();
  const int64_t total = (int64_t)T * Hkv * CPR;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % CPR);
    const int h = (int)((i / CPR) % Hkv);
    const int t = (int)(i / ((int64_t)CPR * Hkv));
    const int64_t slot = slot_mapping[t];
    if (slot < 0) continue;  // PAD_SLOT_ID: token not cached
    const int64_t blk = slot / block_size, off = slot % block_size;
    const uint4 kx = ldg_nc_v4(key + t * key_st + h * key_sh + c * 8);
    const uint4 vx = ldg_nc_v4(value + t * val_st + h * val_sh + c * 8);
    stg_v4(k_cache + blk * kc_sblk + off * kc_stok + h * kc_sh + c * 8, kx);
    stg_v4(v_cache + blk * vc_sblk + off * vc_stok + h * vc_sh + c * 8, vx);
  }
}

__global__ void __launch_bounds__(256)
state_zero_kernel(uint4* __restrict__ state, const int32_t* __restrict__ slots, int64_t st_sl16,
                  int64_t st_ss16, int64_t slot_chunks) {
  const int slot = slots[blockIdx.y];
  if (slot < 0) return;
  uint4* dst = state + blockIdx.z * st_sl16 + (int64_t)slot * st_ss16;
  const uint4 z = make_uint4(0, 0, 0, 0);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < slot_chunks;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = z;
}

__global__ void __launch_bounds__(256)
state_copy_kernel(uint4* __restrict__ state, int from_slot, int to_slot, int64_t st_sl16,
                  int64_t st_ss16, int64_t slot_chunks) {
  const uint4* src = state + blockIdx.z * st_sl16 + (int64_t)from_slot * st_ss16;
  uint4* dst = state + blockIdx.z * st_sl16 + (int64_t)to_slot * st_ss16;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < slot_chunks;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}

// Multi-step decode: advance the device-side metadata by one token per live request (vLLM's
// advance_step_flashattn, called from MetaAttentionMetadata.advance_step, meta_attention.py:338-409).
__global__ void __launch_bounds__(256)
advance_step_kernel(int num_seqs, int num_queries, int block_size, int64_t* __restrict__ input_tokens,
                    const int64_t* __restrict__ sampled_token_ids, int64_t* __restrict__ input_positions,
                    int32_t* __restrict__ seq_lens, int64_t* __restrict__ slot_mapping,
                    const int32_t* __restrict__ block_tables, int64_t bt_stride) {
  pdl_wait();
  pdl_launch_dependents();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < num_seqs; i += gridDim.x * blockDim.x) {
    if (i >= num_queries) {          // CUDA-graph padding rows: nothing is cached for them
      slot_mapping[i] = -1;
      continue;
    }
    if (input_tokens && sampled_token_ids) input_tokens[i] = sampled_token_ids[i];
    const int next_len = seq_lens[i] + 1;
    const int pos = next_len - 1;
    seq_lens[i] = next_len;
    if (input_positions) input_positions[i] = pos;
    const int64_t blk = block_tables[(int64_t)i * bt_stride + pos / block_size];
    slot_mapping[i] = blk * block_size + pos % block_size;
  }
}

}  // namespace hymeta

using namespace hymeta;

extern "C" int hymeta_advance_step(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                                   const int64_t* sampled_token_ids, int64_t* input_positions,
                                   int32_t* seq_lens, int64_t* slot_mapping,
                                   const int32_t* block_tables, int64_t bt_stride, void* stream) {
  if (num_seqs == 0) return HYMETA_OK;
  if (num_seqs < 0 || num_queries < 0 || num_queries > num_seqs || block_size <= 0 || !seq_lens ||
      !slot_mapping || !block_tables || bt_stride < 0)
    return HYMETA_ERR_INVALID_ARG;
  const int blocks = (num_seqs + 255) / 256;
  launch_pdl(advance_step_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, num_seqs,
             num_queries, block_size, input_tokens, sampled_token_ids, input_positions, seq_lens,
             slot_mapping, block_tables, bt_stride);
  count_launch();
  return check_launch();
}

extern "C" int hymeta_kv_write_paged(const void* key, const void* value, void* k_cache,
                                     void* v_cache, const int64_t* slot_mapping, int T, int Hkv,
                                     int D, int block_size, int64_t key_st, int64_t key_sh,
                                     int64_t val_st, int64_t val_sh, int64_t kc_sblk,
                                     int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk,
                                     int64_t vc_stok, int64_t vc_sh, void* stream) {
  if (T == 0) return HYMETA_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping || T < 0 || Hkv <= 0 ||
      block_size <= 0)
    return HYMETA_ERR_INVALID_ARG;
  if (D != kD) return HYMETA_ERR_UNSUPPORTED;
  const int64_t strides[] = {key_st, key_sh, val_st, val_sh, kc_sblk, kc_stok, kc_sh,
                             vc_sblk, vc_stok, vc_sh};
  for (int64_t s : strides)
    if (s % 8) return HYMETA_ERR_INVALID_ARG;  // 16-byte vector access
  const int64_t total = (int64_t)T * Hkv * (kD / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  launch_pdl(kv_write_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream,
      (const __nv_bfloat16*)key, (const __nv_bfloat16*)value, (__nv_bfloat16*)k_cache,
      (__nv_bfloat16*)v_cache, slot_mapping, T, Hkv, block_size, key_st, key_sh, val_st, val_sh,
      kc_sblk, kc_stok, kc_sh, vc_sblk, vc_stok, vc_sh);
  count_launch();
  return check_launch();
}

extern "C" int hymeta_state_zero_slots(void* state, const int32_t* slots, int n, int L,
                                       int64_t st_sl, int64_t st_ss, int64_t slot_elems,
                                       void* stream) {
  if (n == 0 || L == 0) return HYMETA_OK;
  if (!state || !slots || n < 0 || L < 0 || slot_elems <= 0) return HYMETA_ERR_INVALID_ARG;
  if ((st_sl % 8) || (st_ss % 8) || (slot_elems % 8)) return HYMETA_ERR_INVALID_ARG;
  if (n > 65535 || L > 65535) return HYMETA_ERR_UNSUPPORTED;
  const int64_t chunks = slot_elems / 8;
  dim3 grid((unsigned)((chunks + 1023) / 1024 < 64 ? (chunks + 1023) / 1024 : 64), n, L);
  state_zero_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((uint4*)state, slots, st_sl / 8,
                                                            st_ss / 8, chunks);
  count_launch();
  return check_launch();
}

extern "C" int hymeta_state_copy_slot(void* state, int from_slot, int to_slot, int L, int64_t st_sl,
                                      int64_t st_ss, int64_t slot_elems, void* stream) {
  if (L == 0 || from_slot == to_slot) return HYMETA_OK;
  if (!state || from_slot < 0 || to_slot < 0 || L < 0 || slot_elems <= 0)
    return HYMETA_ERR_INVALID_ARG;
  if ((st_sl % 8) || (st_ss % 8) || (slot_elems % 8)) return HYMETA_ERR_INVALID_ARG;
  const int64_t chunks = slot_elems / 8;
  dim3 grid((unsigned)((chunks + 1023) / 1024 < 64 ? (chunks + 1023) / 1024 : 64), 1, L);
  state_copy_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((uint4*)state, from_slot, to_slot,
                                                            st_sl / 8, st_ss / 8, chunks);
  count_launch();
  return check_launch();
}
