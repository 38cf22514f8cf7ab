/*
 * hymeta_b200.h -- C ABI of the B200 (sm_100a) Hymeta hybrid-attention kernels.
 *
 * The reference (FAndromedA/vllm-hymeta) is pure Python and defines no FFI; its
 * kernel boundary is a set of Python functions (Triton / flash-attn / fla
 * wrappers).  Each entry point below replaces one of them and is what the
 * reference-side Python stub binds with ctypes.  Citations
 * are file:line relative to the reference root.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated; the caller owns all memory
 *   - activations / KV / state are bf16 (uint16 storage); gates are fp32 or bf16
 *   - strides are in ELEMENTS; the innermost (head_dim) stride is always 1
 *   - `stream` is a cudaStream_t passed as void*
 *   - no allocation and no synchronisation: every call only enqueues kernels on
 *     `stream` and is CUDA-graph-capture safe.  Process-wide state is limited to
 *     (a) the per-device "max dynamic shared memory" function attribute, set on
 *     the first launch on each device (thread safe), (b) the launch counter and
 *     the thread-local last-error code below, and (c) two tuning knobs read ONCE
 *     from the environment: HYMETA_PDL=0 disables programmatic dependent launch
 *     between the decode-step kernels, HYMETA_SWA_DEC_TPS=<n> overrides the
 *     split-K plan of the SWA decode (tiles per split; 0/unset = built-in rule)
 *   - return 0 on success, a negative hymeta_status otherwise
 *   - head_dim (D) must be 128 (the only geometry the reference models use,
 *     SURVEY F2)
 */
#ifndef HYMETA_B200_H
#define HYMETA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HYMETA_ABI_VERSION 6   /* 3: + rope_kv_write, *_gnorm, branch_*, slot_mgr_*; 4: paged-key prefill,
                                  new_slot_mapping in the fused decode; 5: 7B gate folded into the GLA kernels (f2);
                                  6: balanced single-pass GLA prefill (HYMETA_GLA_SEG_BALANCED), static-scratch decode */

typedef enum {
  HYMETA_OK = 0,
  HYMETA_ERR_INVALID_ARG = -1,   /* null pointer / negative size / bad stride          -> ValueError        */
  HYMETA_ERR_UNSUPPORTED = -2,   /* head size / group size / block size not supported  -> ValueError        */
  HYMETA_ERR_WORKSPACE = -3,     /* workspace too small                                -> ValueError        */
  HYMETA_ERR_CUDA = -4,          /* launch failed; see hymeta_last_cuda_error()        -> RuntimeError      */
  HYMETA_ERR_NO_DEVICE = -5,     /* no sm_100 device / driver entry point missing      -> RuntimeError      */
  HYMETA_ERR_NO_SLOT = -6        /* more live sequences than max_num_seqs state slots  -> RuntimeError      */
} hymeta_status;

int hymeta_abi_version(void);
const char* hymeta_status_string(int status);
/* cudaError_t of the last failed launch on this thread (0 if none). */
int hymeta_last_cuda_error(void);
/* Number of kernels launched by this library in this process (bench.py's gpu_launches). */
int64_t hymeta_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * a2  GLA decode -- replaces my_fused_recurrent_gla / my_fused_recurrent_fwd_kernel
 *     (vllm_hymeta/models/my_fused_recurrent_gla.py:8-23, my_fused_recurrent.py:8-135;
 *      7B twin vllm_hymeta/model_for_7B/my_fused_recurrent.py:8-153).
 *
 * One recurrent step for B requests, state updated IN PLACE:
 *     S <- bf16(k (x) v) + exp(g)[:,None] * S ;  o = (q*scale)^T S   (fp32 math, bf16 stores)
 * q,o: [B, H, D].  k,g: [B, Hk, D].  v: [B, Hv, D].  state: [slots, Hs, D, D] (row = d_k, col = d_v).
 * Hs is the number of states per request: Hs == H reproduces the reference layout
 * (one state per q-head, k/v/g already repeat_kv'ed or not), Hs == Hkv stores one state
 * per kv-head (the 7 q-heads of a GQA group share a bit-identical state -- SURVEY F7).
 * Hk and Hv may differ: with the 70B per-q-head lower bound k and g carry H heads while v stays
 * un-repeated (Hv = Hkv).  Requires H % Hs == 0, Hs % Hk == 0, Hs % Hv == 0: state-head s uses
 * k/g head s / (Hs/Hk) and v head s / (Hs/Hv).  slot_idx[b] == -1 skips the request
 * (my_fused_recurrent.py:31-35); its output row is written as zeros.
 * ---------------------------------------------------------------------------------------------- */
int hymeta_gla_decode(const void* q, const void* k, const void* v, const void* g, void* o,
                      void* state, const int32_t* slot_idx,
                      int B, int H, int Hk, int Hv, int Hs, int D,
                      int64_t q_sb, int64_t q_sh, int64_t k_sb, int64_t k_sh,
                      int64_t v_sb, int64_t v_sh, int64_t g_sb, int64_t g_sh,
                      int64_t o_sb, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                      int g_is_fp32, float scale, void* stream);

/* ------------------------------------------------------------------------------------------------
 * a3  GLA chunked prefill, varlen -- replaces the per-request fused_chunk_gla loop
 *     (vllm_hymeta/models/modeling_hymeta.py:441-549 incl. the meta pass :447-467;
 *      vllm_hymeta/model_for_7B/gla_attention.py:179-221).
 *
 * q,o: [T, H, D].  k,g: [T, Hk, D].  v: [T, Hv, D] (head mapping as in hymeta_gla_decode).
 * Sequence n covers tokens [cu_seqlens[n], cu_seqlens[n+1]).
 * init_mode[n]: 0 = zero state, 1 = state[slot_idx[n]], 2 = meta_state ([Hs, D, D], may be null
 * if unused).  If write_state != 0 and slot_idx[n] >= 0 the final state is stored (bf16) to
 * state[slot_idx[n]].  No host synchronisation (replaces the isnan()/context_len host syncs,
 * modeling_hymeta.py:489,501).  max_seqlen is an upper bound used for grid sizing only.
 * ---------------------------------------------------------------------------------------------- */
int hymeta_gla_prefill_varlen(const void* q, const void* k, const void* v, const void* g, void* o,
                              void* state, const void* meta_state,
                              const int32_t* cu_seqlens, const int32_t* slot_idx,
                              const int32_t* init_mode,
                              int N, int T, int H, int Hk, int Hv, int Hs, int D,
                              int64_t q_st, int64_t q_sh, int64_t k_st, int64_t k_sh,
                              int64_t v_st, int64_t v_sh, int64_t g_st, int64_t g_sh,
                              int64_t o_st, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                              int64_t meta_sh, int g_is_fp32, float scale, int write_state,
                              int max_seqlen, void* stream);

/* Sequence-parallel form of a3 for few long sequences (BASELINE configs 3 and 5): every sequence is
 * cut into segments of seg_len tokens (a multiple of 16) that run as independent CTAs -- (1) each
 * segment's local end state from a zero start plus its total decay, (2) a scan over the segments of
 * each chain, (3) the outputs from the scanned start states.  Same results as the single-pass
 * form up to fp32 re-association of the state.  seg_len == 0 (or max_seqlen <= seg_len) is the
 * single-pass form and needs no workspace.  The fla kernel the reference calls is sequential per
 * (request, head) (modeling_hymeta.py:512-513); the caller owns the fp32 workspace of
 * hymeta_gla_prefill_workspace_bytes(N, Hs, max_seqlen, seg_len) bytes (16-byte aligned).
 * k/g/v strides must be multiples of 8 elements (TMA). */
size_t hymeta_gla_prefill_workspace_bytes(int N, int Hs, int max_seqlen, int seg_len);

/* Balanced single-pass form for MORE chains than SMs (N * Hs > SM count; BASELINE config 2: 256 chains
 * on 148 SMs): pass seg_len = HYMETA_GLA_SEG_BALANCED to the _seg / _seg_gnorm / _fused entry points
 * with a workspace of hymeta_gla_prefill_balanced_workspace_bytes(N, Hs) bytes (16-byte aligned).  A
 * one-CTA plan kernel lays the chains end to end in 16-token steps and cuts the line into one equal
 * range per SM; a chain that straddles a cut hands its fp32 state (and the decay it still owes) from
 * one CTA to the next through the workspace (release / acquire flag), so every SM walks the same
 * number of steps instead of whole chains in waves.  Each token is still read once and the state
 * recurrence is evaluated in the same order with the same roundings: results are bit-identical to
 * the single-pass form.  The launch is one persistent CTA per SM; between two pieces of its list a
 * CTA drains and restarts its pipeline (~4 steps), so the form pays off when chains are long and the
 * last wave of a one-chain-per-CTA grid would be partial (B200: 256 chains x 512 steps -9 %,
 * 160 x 512 -41 %), not for thousands of short chains (4000 x 16 steps +25 %).  When a chain is
 * longer than an SM's share (nothing to balance) or there are no more chains than SMs, whole chains
 * are dealt round-robin. */
#define HYMETA_GLA_SEG_BALANCED (-1)
size_t hymeta_gla_prefill_balanced_workspace_bytes(int N, int Hs);
int hymeta_gla_prefill_varlen_seg(const void* q, const void* k, const void* v, const void* g,
                                  void* o, void* state, const void* meta_state,
                                  const int32_t* cu_seqlens, const int32_t* slot_idx,
                                  const int32_t* init_mode,
                                  int N, int T, int H, int Hk, int Hv, int Hs, int D,
                                  int64_t q_st, int64_t q_sh, int64_t k_st, int64_t k_sh,
                                  int64_t v_st, int64_t v_sh, int64_t g_st, int64_t g_sh,
                                  int64_t o_st, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                                  int64_t meta_sh, int g_is_fp32, float scale, int write_state,
                                  int max_seqlen, int seg_len, void* workspace,
                                  size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * f2  GLA kernels with the 7B per-head g_norm fused as epilogue -- additionally replaces
 *     `o = self.g_norm(o)` (vllm_hymeta/model_for_7B/gla_attention.py:177 RMSNorm(head_dim, eps),
 *     applied at :210 after fused_chunk_gla and :233 after my_fused_recurrent_gla).
 * Same arguments as hymeta_gla_decode / hymeta_gla_prefill_varlen_seg plus gnorm_weight
 * ([D] bf16, 16-byte aligned; null = no norm, identical to the plain entry points) and gnorm_eps.
 * Arithmetic = vLLM's rms_norm op on the bf16 kernel output: x = bf16(o); rstd =
 * rsqrt(mean(x^2) + eps) in fp32; y = bf16(bf16(x * rstd) * w).
 * ---------------------------------------------------------------------------------------------- */
int hymeta_gla_decode_gnorm(const void* q, const void* k, const void* v, const void* g, void* o,
                            void* state, const int32_t* slot_idx, int B, int H, int Hk, int Hv,
                            int Hs, int D, int64_t q_sb, int64_t q_sh, int64_t k_sb, int64_t k_sh,
                            int64_t v_sb, int64_t v_sh, int64_t g_sb, int64_t g_sh,
                            int64_t o_sb, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                            int g_is_fp32, float scale, const void* gnorm_weight, float gnorm_eps,
                            void* stream);
int hymeta_gla_prefill_varlen_seg_gnorm(
    const void* q, const void* k, const void* v, const void* g, void* o, void* state,
    const void* meta_state, const int32_t* cu_seqlens, const int32_t* slot_idx,
    const int32_t* init_mode, int N, int T, int H, int Hk, int Hv, int Hs, int D, int64_t q_st,
    int64_t q_sh, int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t g_st,
    int64_t g_sh, int64_t o_st, int64_t o_sh, int64_t st_ss, int64_t st_sh, int64_t meta_sh,
    int g_is_fp32, float scale, int write_state, int max_seqlen, int seg_len, void* workspace,
    size_t workspace_bytes, const void* gnorm_weight, float gnorm_eps, void* stream);

/* ------------------------------------------------------------------------------------------------
 * a4  GLA feature map + gate pre-processing -- replaces the eager torch elementwise chain between
 *     qkv_proj and the GLA kernels:
 *       70B  vllm_hymeta/models/modeling_hymeta.py:598-629 (relu(q); sigmoid(float(k)); clamp or
 *            lower bound; g = log(.) in fp32; k -> bf16), HF twin hf_model/linear_attention.py:143-157
 *       7B   vllm_hymeta/model_for_7B/gla_attention.py:251-262 (relu(q); relu(k);
 *            g = logsigmoid(gk) / gate_logit_normalizer in bf16)
 * q_in: [T, H, D], k_in (and gk_in): [T, Hk, D] -- strided views of the projection output are fine
 * (strides in elements).  q_out may alias q_in, k_out may alias k_in when Ho == Hk.
 * 70B: lower_bound == null (or the first layer of a group, modeling_hymeta.py:616) -> k_out, g_out
 * have Ho == Hk heads.  lower_bound != null: fp32 [Ho, D], the rank-local slice of the reference's
 * per-q-head row (:620-625); Ho == H reproduces the reference (each q-head its own gate, the raw
 * k-head is broadcast over its GQA group in-kernel, no repeat_kv copy); Ho must be a multiple of Hk.
 * g_out: fp32 [T, Ho, D] (70B) / bf16 [T, Hk, D] (7B).
 * ---------------------------------------------------------------------------------------------- */
int hymeta_gla_gate_70b(const void* q_in, const void* k_in, void* q_out, void* k_out, float* g_out,
                        const float* lower_bound, int T, int H, int Hk, int Ho, int D,
                        int64_t qi_st, int64_t qi_sh, int64_t ki_st, int64_t ki_sh,
                        int64_t qo_st, int64_t qo_sh, int64_t ko_st, int64_t ko_sh,
                        int64_t go_st, int64_t go_sh, float clamp_max, void* stream);
int hymeta_gla_gate_7b(const void* q_in, const void* k_in, const void* gk_in, void* q_out,
                       void* k_out, void* g_out, int T, int H, int Hk, int D,
                       int64_t qi_st, int64_t qi_sh, int64_t ki_st, int64_t ki_sh,
                       int64_t gi_st, int64_t gi_sh, int64_t qo_st, int64_t qo_sh,
                       int64_t ko_st, int64_t ko_sh, int64_t go_st, int64_t go_sh,
                       float normalizer, void* stream);

/* ------------------------------------------------------------------------------------------------
 * f2  7B feature map + gate folded into the GLA kernels (vllm_hymeta/model_for_7B/
 *     gla_attention.py:251-262 and its callers :179-235): the GLA kernels read the RAW qkv_proj /
 *     gk_proj rows, so the 37 KB/token q / k / g round trip of hymeta_gla_gate_7b disappears.
 *
 * hymeta_gla_gate_7b_g:           g = bf16(bf16(logsigmoid(gk)) / normalizer) only (2 KB/token); the
 *                                 prefill kernel takes it together with raw q / k and relu_qk = 1.
 * hymeta_gla_prefill_varlen_fused: hymeta_gla_prefill_varlen_seg_gnorm plus relu_qk (0 / 1): relu is
 *                                 applied to q and k as they are scaled into the MMA operands
 *                                 (relu(x) * e == max(x * e, 0) for e > 0: bit-identical outputs);
 *                                 and gate_table (null, or [65536] bf16): `g` then holds the RAW bf16
 *                                 gk rows and the kernel's gate warps map each through the table --
 *                                 the 7B gate is a pure function of a 16-bit input; the table is
 *                                 hymeta_gla_gate_7b_g applied once to all 65536 bit patterns, so the
 *                                 result is bit-identical to the separate pass and the layer's
 *                                 prefill is ONE launch.  sumsq_out (null, or [T, H] fp32): the O
 *                                 epilogue also emits each bf16 output row's sum of squares (f1: the
 *                                 branch merge's RMS-norm partials).
 * hymeta_gla_decode_gate7b:       hymeta_gla_decode_gnorm plus gate_normalizer: 0 = plain; > 0 = q, k
 *                                 and g are raw q / k / gk rows (g bf16) and relu / logsigmoid /
 *                                 division run in the kernel's operand conversion, bit-identical to
 *                                 hymeta_gla_gate_7b followed by the plain kernel.  gate_normalizer
 *                                 must be a power of two (the model's 16; else HYMETA_ERR_UNSUPPORTED).
 * ---------------------------------------------------------------------------------------------- */
int hymeta_gla_gate_7b_g(const void* gk_in, void* g_out, int T, int Hk, int D, int64_t gi_st,
                         int64_t gi_sh, int64_t go_st, int64_t go_sh, float normalizer, void* stream);
int hymeta_gla_prefill_varlen_fused(
    const void* q, const void* k, const void* v, const void* g, void* o, void* state,
    const void* meta_state, const int32_t* cu_seqlens, const int32_t* slot_idx,
    const int32_t* init_mode, int N, int T, int H, int Hk, int Hv, int Hs, int D, int64_t q_st,
    int64_t q_sh, int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t g_st,
    int64_t g_sh, int64_t o_st, int64_t o_sh, int64_t st_ss, int64_t st_sh, int64_t meta_sh,
    int g_is_fp32, float scale, int write_state, int max_seqlen, int seg_len, void* workspace,
    size_t workspace_bytes, const void* gnorm_weight, float gnorm_eps, int relu_qk,
    const void* gate_table, float* sumsq_out, void* stream);
int hymeta_gla_decode_gate7b(const void* q, const void* k, const void* v, const void* g, void* o,
                             void* state, const int32_t* slot_idx, int B, int H, int Hk, int Hv,
                             int Hs, int D, int64_t q_sb, int64_t q_sh, int64_t k_sb, int64_t k_sh,
                             int64_t v_sb, int64_t v_sh, int64_t g_sb, int64_t g_sh,
                             int64_t o_sb, int64_t o_sh, int64_t st_ss, int64_t st_sh,
                             int g_is_fp32, float scale, const void* gnorm_weight, float gnorm_eps,
                             float gate_normalizer, void* stream);

/* ------------------------------------------------------------------------------------------------
 * a9  Paged-KV write -- replaces torch.ops._C_cache_ops.reshape_and_cache_flash
 *     (vllm_hymeta/attention/meta_attention.py:790-799).  Bit-exact copy:
 *     cache[slot / block_size, slot % block_size, h, :] = key|value[t, h, :]   for slot >= 0.
 * key,value: [T, Hkv, D].  k_cache,v_cache: [num_blocks, block_size, Hkv, D].
 * ---------------------------------------------------------------------------------------------- */
int hymeta_kv_write_paged(const void* key, const void* value, void* k_cache, void* v_cache,
                          const int64_t* slot_mapping, int T, int Hkv, int D, int block_size,
                          int64_t key_st, int64_t key_sh, int64_t val_st, int64_t val_sh,
                          int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh,
                          int64_t vc_sblk, int64_t vc_stok, int64_t vc_sh, void* stream);

/* ------------------------------------------------------------------------------------------------
 * f3  Fused SWA prologue: NeoX rotary embedding of q and k + paged-KV write + meta-K/V capture --
 *     replaces, in one launch, HymetaRotaryEmbedding.forward -> ops.rotary_embedding
 *     (vllm_hymeta/models/modeling_hymeta.py:220-234, called at :762), the position shift and
 *     meta-position concat (:1314-1324), reshape_and_cache_flash (meta_attention.py:790-799)
 *     and the two meta_fattn_cache copy_ calls (modeling_hymeta.py:776-792).
 * q_in [T,H,D], k_in,v_in [T,Hkv,D] (views of the qkv_proj output); q_out,k_out may alias the
 * inputs (the reference op is in place).  cos_sin_cache: the module's bf16 table [max_pos, D]
 * = [cos(pos*inv_freq) | sin(pos*inv_freq)] (modeling_hymeta.py:210-218, cache_dtype bf16 :1187).
 * Arithmetic, fp32_math == 0: the rotary op of the vLLM the reference pins (0.10.0,
 * csrc/pos_encoding_kernels.cu instantiated for c10::BFloat16): each product and the add/sub
 * rounded to bf16 (bit-exact against torch bf16 arithmetic; that vLLM is not installable here, so
 * this mode is pinned to the source text only).  fp32_math != 0: products and add/sub in fp32, ONE
 * rounding to bf16 -- bit-identical to torch.ops._C.rotary_embedding of the vLLM 0.22 installed
 * next to this package, checked live on B200 (tests/test_gpu_live_libs.py).
 * Tokens t < n_meta are the meta tokens of the first pass: position t, k/v rows stored into
 * meta_k/meta_v [n_meta,Hkv,D] instead of pages.  Token t >= n_meta: position
 * positions[t-n_meta] + pos_offset (pos_offset = num_meta_tokens, :1314), rotated k and v stored
 * at slot_mapping[t-n_meta] (< 0: not cached).  k_cache == null (profiling run): rotate only.
 * ---------------------------------------------------------------------------------------------- */
int hymeta_rope_kv_write_paged(
    const void* q_in, const void* k_in, const void* v_in, void* q_out, void* k_out,
    const void* cos_sin_cache, int max_pos, const int64_t* positions, int64_t pos_offset,
    void* k_cache, void* v_cache, const int64_t* slot_mapping, void* meta_k, void* meta_v,
    int n_meta, int T, int H, int Hkv, int D, int block_size,
    int64_t qi_st, int64_t qi_sh, int64_t ki_st, int64_t ki_sh, int64_t vi_st, int64_t vi_sh,
    int64_t qo_st, int64_t qo_sh, int64_t ko_st, int64_t ko_sh,
    int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh, int64_t vc_sblk, int64_t vc_stok,
    int64_t vc_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, int fp32_math,
    void* stream);

/* ------------------------------------------------------------------------------------------------
 * a8  SWA + meta-prefix paged decode -- replaces metatoken_flash_attn_with_kvcache
 *     (vllm_hymeta/attention/fa_kernel_metatoken.py:287-326 -> :124-194 -> :26-106,
 *      i.e. flash_attn fwd_kvcache + meta pass + _update_out_and_lse in ONE pass).
 *
 * q,out: [B, H, D].  Request b attends to [meta_k/meta_v (M rows, always visible) |
 * cached keys at logical positions max(0, L-1-window_left) .. L-1], L = cache_seqlens[b]
 * (window_left < 0: all positions).  Logical position p lives at
 * cache[block_tables[b, p / block_size], p % block_size].  meta_k/meta_v: [M, Hkv, D] or null
 * when M == 0.  num_blocks = physical blocks in the cache (extent of the TMA tensor map),
 * max_blocks = columns of block_tables.  block_size must be a multiple of 16 (as the reference
 * requires, meta_attention.py:88-91).  workspace: hymeta_swa_decode_workspace_bytes(...) bytes,
 * any contents.
 * ---------------------------------------------------------------------------------------------- */
size_t hymeta_swa_decode_workspace_bytes(int B, int H, int Hkv, int D, int max_blocks,
                                         int block_size, int window_left, int M);
int hymeta_swa_meta_decode_paged(const void* q, const void* k_cache, const void* v_cache,
                                 const int32_t* block_tables, const int32_t* cache_seqlens,
                                 const void* meta_k, const void* meta_v, void* out,
                                 void* workspace, size_t workspace_bytes,
                                 int B, int H, int Hkv, int D, int M, int block_size,
                                 int num_blocks, int max_blocks, int64_t bt_stride,
                                 int window_left, float scale,
                                 int64_t q_sb, int64_t q_sh, int64_t o_sb, int64_t o_sh,
                                 int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh,
                                 int64_t vc_sblk, int64_t vc_stok, int64_t vc_sh,
                                 int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
                                 void* stream);

/* a8 + a9 fused: the same decode with (1) the step's paged-KV write folded in --
 * replaces reshape_and_cache_flash (meta_attention.py:790-799) followed by the decode call
 * (:954-981): new_k/new_v [B, Hkv, D] are the K/V of the token at position cache_seqlens[b]-1; the
 * kernel stores them to their page and attends to them from registers, so the pages need not hold
 * that row yet (null: the pages already hold it).  The destination is new_slot_mapping[b] in
 * reshape_and_cache_flash's convention (slot = block * block_size + offset; slot < 0 = padded row,
 * nothing is stored -- CUDA-graph padding, meta_attention.py:557-562), or, when new_slot_mapping
 * is null, cache[block_tables[b, p / block_size], p % block_size] with p = cache_seqlens[b]-1 (then
 * padded rows MUST carry cache_seqlens[b] = 0 so that no store happens) -- and
 * (2) the split-K combine folded into the split kernel: `tickets` is
 * hymeta_swa_decode_ticket_bytes(B, H, Hkv) bytes of int32, ZERO before the first call; the kernel
 * leaves it zero (null: separate combine launch).  One launch per layer-step instead of three. */
size_t hymeta_swa_decode_ticket_bytes(int B, int H, int Hkv);
int hymeta_swa_meta_decode_paged_fused(const void* q, void* k_cache, void* v_cache,
                                       const int32_t* block_tables, const int32_t* cache_seqlens,
                                       const void* meta_k, const void* meta_v, void* out,
                                       void* workspace, size_t workspace_bytes,
                                       int B, int H, int Hkv, int D, int M, int block_size,
                                       int num_blocks, int max_blocks, int64_t bt_stride,
                                       int window_left, float scale,
                                       int64_t q_sb, int64_t q_sh, int64_t o_sb, int64_t o_sh,
                                       int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh,
                                       int64_t vc_sblk, int64_t vc_stok, int64_t vc_sh,
                                       int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
                                       const void* new_k, const void* new_v,
                                       int64_t nk_sb, int64_t nk_sh, int64_t nv_sb, int64_t nv_sh,
                                       const int64_t* new_slot_mapping, int32_t* tickets,
                                       void* stream);

/* No-allocation variant of hymeta_swa_meta_decode_paged.  It has NO counterpart in the reference: it
 * was added for callers that cannot call cudaMalloc (code
 * under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same arguments and
 * results as hymeta_swa_meta_decode_paged minus the workspace: the split-K partials and the combine
 * tickets live in a static device array inside the library (HYMETA_SWA_DECODE_STATIC_WS_BYTES +
 * HYMETA_SWA_DECODE_STATIC_TICKET_BYTES bytes on every GPU that loads the library, held until
 * exit), and the call performs no cudaMalloc / cudaFree / allocating library call -- one
 * cudaMemsetAsync of the tickets and one kernel launch on `stream`.  Limits: a plan whose partials
 * (hymeta_swa_decode_workspace_bytes) exceed HYMETA_SWA_DECODE_STATIC_WS_BYTES or whose tickets
 * (hymeta_swa_decode_ticket_bytes) exceed HYMETA_SWA_DECODE_STATIC_TICKET_BYTES returns
 * HYMETA_ERR_WORKSPACE and launches nothing (16 MiB covers B x H x splits <= 32 512 rows, e.g.
 * B = 1000 x 28 heads in one split, or B = 64 x 28 heads in up to 18 splits).  All calls on one GPU
 * share the array: calls that may overlap on the GPU (different streams without an ordering
 * between them) race on it -- serialise them, or use the caller-workspace entry points. */
#define HYMETA_SWA_DECODE_STATIC_WS_BYTES (16u << 20)
#define HYMETA_SWA_DECODE_STATIC_TICKET_BYTES (64u << 10)
int hymeta_swa_meta_decode_paged_static(const void* q, const void* k_cache, const void* v_cache,
                                        const int32_t* block_tables, const int32_t* cache_seqlens,
                                        const void* meta_k, const void* meta_v, void* out,
                                        int B, int H, int Hkv, int D, int M, int block_size,
                                        int num_blocks, int max_blocks, int64_t bt_stride,
                                        int window_left, float scale,
                                        int64_t q_sb, int64_t q_sh, int64_t o_sb, int64_t o_sh,
                                        int64_t kc_sblk, int64_t kc_stok, int64_t kc_sh,
                                        int64_t vc_sblk, int64_t vc_stok, int64_t vc_sh,
                                        int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
                                        void* stream);

/* ------------------------------------------------------------------------------------------------
 * a10 multi-step decode metadata advance -- replaces vllm._custom_ops.advance_step_flashattn as
 *     called by MetaAttentionMetadata.advance_step (vllm_hymeta/attention/meta_attention.py:338-409).
 *
 * For i < num_queries: seq_lens[i] += 1; p = seq_lens[i] - 1; input_positions[i] = p;
 * input_tokens[i] = sampled_token_ids[i]; slot_mapping[i] = block_tables[i, p / block_size] *
 * block_size + p % block_size.  For num_queries <= i < num_seqs (CUDA-graph padding):
 * slot_mapping[i] = -1.  input_tokens / sampled_token_ids / input_positions may be null (the
 * attention path needs only seq_lens and slot_mapping).  Exact integer work.
 * ---------------------------------------------------------------------------------------------- */
int hymeta_advance_step(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                        const int64_t* sampled_token_ids, int64_t* input_positions,
                        int32_t* seq_lens, int64_t* slot_mapping, const int32_t* block_tables,
                        int64_t bt_stride, void* stream);

/* ------------------------------------------------------------------------------------------------
 * a7  SWA + meta-prefix varlen prefill -- replaces metatoken_flash_attn_varlen_func
 *     (vllm_hymeta/attention/fa_kernel_metatoken.py:329-374 -> :198-285: two flash-attn
 *      launches + TorchScript LSE merge + concat, in ONE kernel).
 *
 * q1: [Tq, H, D] packed queries, k1,v1: [Tk, Hkv, D] packed keys; sequence n has queries
 * [cu_q[n], cu_q[n+1]) and keys [cu_k[n], cu_k[n+1]); queries are aligned to the END of the keys
 * (flash-attn bottom-right causal).  Real query at key-position i sees the M meta keys plus real
 * keys j <= i with i - j <= window_left (window_left < 0: unbounded).
 * q2: [M, H, D] meta queries or null; when given they attend to the meta keys only,
 * bidirectionally (:261-275), and their rows come FIRST in `out` (:283):
 * out: [(M if q2) + Tq, H, D].
 * If block_tables != null the keys are read from the PAGED cache instead -- prefix-cache hit /
 * chunked prefill (meta_attention.py:886-919) and multi-token decode (:924-953): k1,v1 are then the
 * [num_blocks, block_size, Hkv, D] caches (k_st/v_st = row stride inside a block, kc_sblk/vc_sblk =
 * block stride, Tk = num_blocks * block_size, block_size % 16 == 0), block_tables is
 * [N, max_blocks] with row stride bt_stride, and cu_k holds cumulative seq_lens: request n owns
 * logical key positions 0 .. cu_k[n+1]-cu_k[n]-1, position p at
 * cache[block_tables[n, p / block_size], p % block_size]; its queries are the LAST
 * cu_q[n+1]-cu_q[n] positions.  Cache rows outside the visible range are never used as data (they
 * may hold recycled-page garbage / NaN).  The reference passes the fresh key tensor together with
 * block_tables here, which flash-attn's paged API cannot accept; this implements the
 * evident intent (keys through the block table), like vLLM's stock backend.
 * ---------------------------------------------------------------------------------------------- */
int hymeta_swa_meta_prefill_varlen(const void* q1, const void* k1, const void* v1, const void* q2,
                                   const void* meta_k, const void* meta_v, void* out,
                                   const int32_t* cu_q, const int32_t* cu_k,
                                   const int32_t* block_tables, int64_t bt_stride, int block_size,
                                   int N, int Tq, int Tk, int H, int Hkv, int D, int M,
                                   int max_seqlen_q, int window_left, float scale,
                                   int64_t q_st, int64_t q_sh, int64_t k_st, int64_t k_sh,
                                   int64_t v_st, int64_t v_sh, int64_t q2_st, int64_t q2_sh,
                                   int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh,
                                   int64_t o_st, int64_t o_sh, int64_t kc_sblk, int64_t vc_sblk,
                                   void* stream);

/* ------------------------------------------------------------------------------------------------
 * a5  Slot-state maintenance -- replaces the advanced-index zero fill
 *     `hymeta_cache[:, slots, ...] = 0` (modeling_hymeta.py:1202-1227) and
 *     HymetaCacheManager._copy_cache (models/hymeta_cache.py:46-50) with one launch each.
 * state: [L, slots, Hs, D, D] bf16, contiguous per (layer, slot).
 * ---------------------------------------------------------------------------------------------- */
int hymeta_state_zero_slots(void* state, const int32_t* slots, int n, int L,
                            int64_t st_sl, int64_t st_ss, int64_t slot_elems, void* stream);
int hymeta_state_copy_slot(void* state, int from_slot, int to_slot, int L,
                           int64_t st_sl, int64_t st_ss, int64_t slot_elems, void* stream);

/* ------------------------------------------------------------------------------------------------
 * f1  Branch merge of the 70B hybrid layer -- replaces norm1(out_attn), norm2(out_linear),
 *     (a + b) / 2 and .to(bfloat16) (vllm_hymeta/models/modeling_hymeta.py:909-919) with
 *     HymetaRMSNormTP._forward (:151-173) for both branches.
 * x_attn, x_linear: [T, hidden_local] bf16 (hidden_local = hidden_size / tp_world, the rank's
 * slice).  hymeta_branch_sumsq writes means_out[0][t], means_out[1][t] = mean over the local
 * slice of x^2 (fp32, [2][T]).  With tp_world > 1 the caller all-reduces (SUM) that 8*T-byte
 * buffer across the tensor-parallel group -- the one exchange adjacent to the path (:159-161) --
 * and passes it as `means`; with tp_world == 1 `means` may be null and the kernel computes it.
 * hymeta_branch_norm_merge: out[t] = bf16( ( bf16(x_attn*rstd_a) * w_attn
 *                                          + bf16(x_linear*rstd_l) * w_linear ) / 2 ),
 * rstd = rsqrt(means / tp_world + eps), w_* = the rank's fp32 weight slice [hidden_local].
 * `out` [T, hidden_local] bf16 may alias either input.
 * ---------------------------------------------------------------------------------------------- */
int hymeta_branch_sumsq(const void* x_attn, const void* x_linear, float* means_out, int T,
                        int hidden_local, int64_t xa_st, int64_t xl_st, void* stream);
int hymeta_branch_norm_merge(const void* x_attn, const void* x_linear, const float* means,
                             const float* w_attn, const float* w_linear, void* out, int T,
                             int hidden_local, int tp_world, float eps, int64_t xa_st,
                             int64_t xl_st, int64_t o_st, void* stream);

/* f1, tensor-parallel form in ONE kernel (compute -> exchange over NVLink peer memory -> merge; replaces
 * hymeta_branch_sumsq + the caller's all-reduce + hymeta_branch_norm_merge at tp_world >= 2):
 * every rank owns an exchange buffer of hymeta_branch_merge_tp_buffer_bytes(tp_world, t_cap) bytes,
 * zero-filled once, mapped into every peer (CUDA IPC / symmetric memory); peer_buffers is a HOST array
 * of the tp_world device addresses, indexed by rank, as seen from this process.  A warp reduces a row,
 * stores {mean_attn, epoch, mean_linear, epoch} into each peer's buffer with one 16-byte store, polls
 * its own buffer for the peers' slots of that row, sums the means in rank order and finishes the row
 * (its second touch is an L2 hit: x is read from DRAM once).  epoch: any non-zero value that changes by one between consecutive
 * calls on the group (two slot parities); every rank passes the same T, hidden_local and epoch.  All
 * CTAs are co-resident (grid <= SMs x occupancy).  A peer that never arrives traps after ~1-2 minutes of polling. */
size_t hymeta_branch_merge_tp_buffer_bytes(int tp_world, int64_t t_cap);
int hymeta_branch_norm_merge_tp(const void* x_attn, const void* x_linear, const float* w_attn,
                                const float* w_linear, void* out, int T, int hidden_local,
                                int tp_world, int tp_rank, const uint64_t* peer_buffers,
                                int64_t t_cap, uint32_t epoch, float eps, int64_t xa_st,
                                int64_t xl_st, int64_t o_st, void* stream);
/* Same, with the per-(token, head) sums of squares the attention kernels' epilogues emitted
 * (hymeta_swa_meta_prefill_varlen_sumsq / hymeta_gla_prefill_varlen_fused: [T, heads_local] fp32, row
 * stride heads_local): the merge then reads x exactly once -- SURVEY 8f f1 as specified.  Either both
 * or neither of sumsq_attn / sumsq_linear. */
int hymeta_branch_norm_merge_tp_sumsq(const void* x_attn, const void* x_linear, const float* w_attn,
                                      const float* w_linear, void* out, int T, int hidden_local,
                                      int tp_world, int tp_rank, const uint64_t* peer_buffers,
                                      int64_t t_cap, uint32_t epoch, float eps, int64_t xa_st,
                                      int64_t xl_st, int64_t o_st, const float* sumsq_attn,
                                      const float* sumsq_linear, int heads_local, void* stream);
/* CUDA-graph form of hymeta_branch_norm_merge_tp_sumsq: a replayed graph cannot carry a fresh by-value
 * epoch, so the epoch lives in device memory.  epoch_state: two uint32 {last epoch, CTAs done}, 8-byte
 * aligned, zero before the first call, owned by the exchange (one per peer_buffers set) and touched by
 * nothing else; every CTA derives the call's epoch from it and the last CTA to finish advances it.
 * Every rank must issue (or replay) the same sequence of calls on the group; do not mix this entry
 * point and the by-value one on the same buffers. */
int hymeta_branch_norm_merge_tp_graph(const void* x_attn, const void* x_linear, const float* w_attn,
                                      const float* w_linear, void* out, int T, int hidden_local,
                                      int tp_world, int tp_rank, const uint64_t* peer_buffers,
                                      int64_t t_cap, uint32_t* epoch_state, float eps, int64_t xa_st,
                                      int64_t xl_st, int64_t o_st, const float* sumsq_attn,
                                      const float* sumsq_linear, int heads_local, void* stream);
/* hymeta_swa_meta_prefill_varlen plus sumsq_out ([rows, H] fp32, rows as `out`; null = plain): each
 * output row's sum of squares over the head's 128 bf16 values, from the kernel's epilogue. */
int hymeta_swa_meta_prefill_varlen_sumsq(
    const void* q1, const void* k1, const void* v1, const void* q2, const void* meta_k,
    const void* meta_v, void* out, const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
    const int32_t* block_tables, int64_t bt_stride, int block_size, int N, int Tq, int Tk, int H,
    int Hkv, int D, int M, int max_seqlen_q, int window_left, float scale, int64_t q_st,
    int64_t q_sh, int64_t k_st, int64_t k_sh, int64_t v_st, int64_t v_sh, int64_t q2_st,
    int64_t q2_sh, int64_t mk_st, int64_t mk_sh, int64_t mv_st, int64_t mv_sh, int64_t o_st,
    int64_t o_sh, int64_t kc_sblk, int64_t vc_sblk, float* sumsq_out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * f4  Host-side state-slot manager -- replaces the bookkeeping of vLLM 0.10.0's ConstantSizeCache
 *     inherited by HymetaCacheManager / GLACacheManager (vllm_hymeta/models/hymeta_cache.py:24-63,
 *     vllm_hymeta/model_for_7B/gla_cache.py:17-32; mapping read at modeling_hymeta.py:1206-1208)
 *     and batches the per-request slot clearing of modeling_hymeta.py:1202-1227.  No CUDA calls.
 * prepare(): releases `finished`, then resolves the n (request id, sequence id) pairs of the run
 * in batch order into slots_out (-1 = PAD_SLOT_ID for a request finished before it ran).  A new
 * request pops the BACK of the free list (reference order); its slot is appended to new_slots
 * (zero these with hymeta_state_zero_slots).  A new sequence of a live request forks the
 * request's first sequence: (copy_from[i], copy_to[i]) pairs to pass to hymeta_state_copy_slot, in
 * order.  copy_from/copy_to/new_slots need room for n entries.  HYMETA_ERR_NO_SLOT when the free
 * list runs dry (earlier pairs of the call stay assigned, as in the reference).
 * slot_of(): slot or -1.  free_list(): copies up to cap entries, returns the full count.
 * Id lists (req_ids, finished) are PACKED: n NUL-terminated strings back to back.
 * Not thread-safe per manager (one Python thread per worker).
 * ---------------------------------------------------------------------------------------------- */
typedef struct hymeta_slot_mgr hymeta_slot_mgr;
hymeta_slot_mgr* hymeta_slot_mgr_create(int max_slots);
void hymeta_slot_mgr_destroy(hymeta_slot_mgr* mgr);
int hymeta_slot_mgr_release(hymeta_slot_mgr* mgr, const char* finished, int n_finished);
int hymeta_slot_mgr_prepare(hymeta_slot_mgr* mgr, const char* req_ids,
                            const int64_t* seq_ids, int n, const char* finished,
                            int n_finished, int32_t* slots_out, int32_t* copy_from,
                            int32_t* copy_to, int* n_copies, int32_t* new_slots, int* n_new);
int hymeta_slot_mgr_free_count(const hymeta_slot_mgr* mgr);
int hymeta_slot_mgr_free_list(const hymeta_slot_mgr* mgr, int32_t* out, int cap);
int hymeta_slot_mgr_slot_of(const hymeta_slot_mgr* mgr, const char* req_id, int64_t seq_id);
int hymeta_slot_mgr_live_count(const hymeta_slot_mgr* mgr);
/* Enumerate live (request id, sequence id, slot) triples; ids NUL-separated, one per triple.
 * Returns the triple count, *ids_bytes = bytes the ids need; writes only within the capacities. */
int hymeta_slot_mgr_live(const hymeta_slot_mgr* mgr, char* ids_buf, size_t ids_cap,
                         size_t* ids_bytes, int64_t* seq_ids, int32_t* slots, int cap);

/* Built-in self test of the tcgen05 / TMA building blocks (descriptor encodings, swizzles).
 * Runs small GEMMs on the current device and compares with a CUDA-core result.
 * Returns 0 if all pass, else the (positive) index of the first failing case. */
int hymeta_selftest_umma(float* max_err_out /* host pointer, may be null */);

#ifdef __cplusplus
}
#endif
#endif /* HYMETA_B200_H */
