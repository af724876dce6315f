/*
 * aphrodite_b200.h - C ABI of the B200-native decode hot path.
 *
 * The reference has no C ABI: its boundary is "C++ functions taking torch::Tensor&, registered
 * by schema string" (/root/reference/kernels/torch_bindings.cpp:41-65,326-333,
 * 696-702). This header is the torch-free layer underneath that boundary: plain pointers, sizes
 * and a CUDA stream. aphrodite-fork_b200/csrc/torch_binding.cpp registers the reference's four
 * schema strings on top of it, so `torch.ops._C.*` / `torch.ops._C_cache_ops.*` callers
 * (aphrodite/_custom_ops.py) are unchanged, and any other host (ctypes, cgo, JNI) can bind the
 * same entry points directly.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - all calls are asynchronous on `stream`, never synchronise, never allocate: scratch is
 *     passed in by the caller (sizes from the *_workspace_bytes functions) so that the calls
 *     are CUDA-graph capturable, like the reference ops (worker/model_runner.py:1487-1646);
 *   - programmatic dependent launch (PDL): every kernel is launched with the programmatic-stream-serialisation
 *     attribute and reads what it treats as IMMUTABLE - GEMM weights / scales / zero points, norm weights, the
 *     permutation of aph_permute_cols - BEFORE it waits for the previous kernel of the stream. Contract: those tensors
 *     must not be written by a kernel that triggers its dependents early (i.e. another PDL kernel of this library) on
 *     the same stream right before the call; this library's own weight-producing kernels (aph_*_marlin_repack,
 *     aph_convert_fp8) are plain launches for that reason. Activations, KV caches, block tables and all outputs are
 *     only touched after the wait. APH_PDL=0 turns the attribute off;
 *   - return value: 0 on success, an aph_status_t otherwise; aph_b200_last_error() returns a
 *     thread-local message (the torch binding turns it into c10::Error / RuntimeError, the
 *     reference's error behaviour: TORCH_CHECK, SURVEY.md 8b "Errors").
 */
#ifndef APHRODITE_B200_H_
#define APHRODITE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define APH_B200_ABI_VERSION 1

typedef enum {
  APH_OK = 0,
  APH_ERR_INVALID_ARGUMENT = 1, /* shape / alignment / value the reference would TORCH_CHECK */
  APH_ERR_UNSUPPORTED = 2,      /* legal in the reference, not implemented here yet          */
  APH_ERR_WORKSPACE = 3,        /* scratch buffer too small                                   */
  APH_ERR_CUDA = 4              /* launch failed (cudaGetLastError)                           */
} aph_status_t;

/* scalar_t of query / key / value / a (reference: at::ScalarType Float / Half / BFloat16) */
typedef enum { APH_DT_F32 = 0, APH_DT_F16 = 1, APH_DT_BF16 = 2 } aph_dtype_t;

/* reference: enum Fp8KVCacheDataType, kernels/attention/dtype_fp8.cuh:14-18 */
typedef enum { APH_KV_AUTO = 0, APH_KV_FP8_E4M3 = 1, APH_KV_FP8_E5M2 = 2 } aph_kv_dtype_t;

int aph_b200_abi_version(void);
const char* aph_b200_last_error(void);

/* "auto" | "fp8" | "fp8_e4m3" | "fp8_e5m2" -> aph_kv_dtype_t, -1 if unknown
 * (reference: DISPATCH_BY_KV_CACHE_DTYPE, kernels/quantization/fp8/nvidia/quant_utils.cuh:532-570) */
int aph_kv_dtype_from_string(const char* kv_cache_dtype);

/* ---------------------------------------------------------------------------------------------
 * reshape_and_cache - replaces kernels/cache_kernels.cu:369-395 (kernel :210-262).
 *   key, value   [num_tokens, num_kv_heads, head_size] scalar dtype, row strides in ELEMENTS
 *   key_cache    [num_blocks, num_kv_heads, head_size/x, block_size, x]
 *   value_cache  [num_blocks, num_kv_heads, head_size, block_size]
 *                cache element = scalar dtype (AUTO) or uint8 (FP8); x = 16 / sizeof(cache elt)
 *   slot_mapping [num_tokens] int64, slot < 0 = padding token, skipped
 *   k_scale, v_scale: device pointers to one float each (only read for FP8)
 * Bit-exact with the reference, including the FP8 bytes.
 * ------------------------------------------------------------------------------------------- */
int aph_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                          const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                          int64_t value_stride, int num_kv_heads, int head_size, int block_size, int x,
                          int64_t cache_block_stride, int dtype, int kv_dtype, const float* k_scale,
                          const float* v_scale, void* stream);

/* convert_fp8 - replaces kernels/cache_kernels.cu:499-577 (also accepts E5M2, which the
 * reference rejects). dst/src: one of them is uint8 (fp8), the other `dtype`; n elements.
 * direction: 0 = scalar -> fp8 (dst = fp8(src / scale)), 1 = fp8 -> scalar (dst = src * scale). */
int aph_convert_fp8(void* dst, const void* src, int64_t n, float scale, int dtype, int kv_dtype,
                    int direction, void* stream);

/* ---------------------------------------------------------------------------------------------
 * The rest of kernels/cache_kernels.cu that a block manager calls.
 *
 * reshape_and_cache_flash - replaces cache_kernels.cu:412-497 (kernel :264-306). Caches in the NHD layout
 *   key_cache, value_cache [num_blocks, block_size, num_heads, head_size]; block_stride / page_stride / head_stride
 *   are the cache strides of dims 0 / 1 / 2 in ELEMENTS (the caller may pass a padded or sliced cache).
 *   Bit-exact with the reference, including the FP8 bytes.
 * copy_blocks - replaces cache_kernels.cu:113-170 (kernel :71-93). key_cache_ptrs / value_cache_ptrs: DEVICE arrays of
 *   num_layers cache base addresses; block_mapping: DEVICE int64 [num_pairs, 2] (src block, dst block);
 *   block_bytes = bytes of one block (stride of dim 0 in bytes). Copies both caches of every layer.
 * swap_blocks - replaces cache_kernels.cu:25-67. block_mapping_host: HOST int64 [num_pairs, 2]; kind 0 = device to
 *   device, 1 = device to host, 2 = host to device; asynchronous copies on `stream`, consecutive pairs merged.
 * ------------------------------------------------------------------------------------------- */
int aph_reshape_and_cache_flash(const void* key, const void* value, void* key_cache, void* value_cache,
                                const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                                int64_t value_stride, int64_t block_stride, int64_t page_stride, int64_t head_stride,
                                int num_heads, int head_size, int block_size, int dtype, int kv_dtype,
                                const float* k_scale, const float* v_scale, void* stream);
int aph_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs, const int64_t* block_mapping,
                    int num_layers, int64_t num_pairs, int64_t block_bytes, void* stream);
int aph_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host, int64_t num_pairs,
                    int64_t block_bytes, int kind, void* stream);

/* ---------------------------------------------------------------------------------------------
 * paged attention - replaces kernels/attention/paged_attention_v1.cu:170-192 and
 * paged_attention_v2.cu:176-201 (kernel attention_kernels.cuh:86-495, reduce :563-668).
 * One implementation serves both entry points: a GQA-packed split-K kernel whose partition
 * size is chosen from the problem size (not the API's 512), plus a reduce kernel.
 *   out          [num_seqs, num_heads, head_size] scalar dtype, row stride out_stride (elements)
 *   query        [num_seqs, num_heads, head_size] scalar dtype, row stride q_stride (elements)
 *   key_cache / value_cache as above; kv_block_stride / kv_head_stride in ELEMENTS
 *   block_tables [num_seqs, max_num_blocks_per_seq] int32; seq_lens [num_seqs] int32
 *   alibi_slopes [num_heads] float or NULL
 *   max_seq_len  static upper bound on seq_lens (CUDA-graph rule)
 *   workspace    >= aph_paged_attention_workspace_bytes(...) bytes, 16-byte aligned
 * v2's caller-allocated exp_sums / max_logits / tmp_out are accepted and left untouched (they
 * are sized for 512-token partitions and never read back by the caller, paged_attn.py:155-165).
 * ------------------------------------------------------------------------------------------- */
size_t aph_paged_attention_workspace_bytes(int num_seqs, int num_heads, int num_kv_heads, int head_size,
                                           int max_seq_len);

int aph_paged_attention_v1(void* out, const void* query, const void* key_cache, const void* value_cache,
                           int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                           const int32_t* block_tables, const int32_t* seq_lens, int block_size,
                           int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes,
                           int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                           int64_t kv_head_stride, int dtype, int kv_dtype, const float* k_scale,
                           const float* v_scale, int tp_rank, int blocksparse_local_blocks,
                           int blocksparse_vert_stride, int blocksparse_block_size,
                           int blocksparse_head_sliding_step, void* workspace, size_t workspace_bytes,
                           void* stream);

int aph_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
                           const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
                           int num_kv_heads, int head_size, float scale, const int32_t* block_tables,
                           const int32_t* seq_lens, int block_size, int max_seq_len,
                           int max_num_blocks_per_seq, const float* alibi_slopes, int64_t q_stride,
                           int64_t out_stride, int64_t kv_block_stride, int64_t kv_head_stride, int dtype,
                           int kv_dtype, const float* k_scale, const float* v_scale, int tp_rank,
                           int blocksparse_local_blocks, int blocksparse_vert_stride,
                           int blocksparse_block_size, int blocksparse_head_sliding_step, void* workspace,
                           size_t workspace_bytes, void* stream);

/* aph_paged_attention_v1_noalloc - NO counterpart in the reference; added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same arguments and result as aph_paged_attention_v1 without the workspace: the split-K partials go to a file-scope
 * `static __device__` array of APH_PA_NOALLOC_MAX_SCRATCH_BYTES = 16 MiB inside the library (held by every
 * GPU that loads the library until the process exits; its device address is resolved once per device with
 * cudaGetSymbolAddress on the first call made with that device current). The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine). Problems whose aph_paged_attention_workspace_bytes(...) exceed the 16 MiB are
 * rejected with APH_ERR_WORKSPACE. All calls on a device share the one array: calls that overlap on the GPU (two
 * streams, or two host threads on different streams) race on it - serialise them on one stream. */
#define APH_PA_NOALLOC_MAX_SCRATCH_BYTES (16u << 20)
int aph_paged_attention_v1_noalloc(void* out, const void* query, const void* key_cache, const void* value_cache,
                                   int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                                   const int32_t* block_tables, const int32_t* seq_lens, int block_size,
                                   int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes,
                                   int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                                   int64_t kv_head_stride, int dtype, int kv_dtype, const float* k_scale,
                                   const float* v_scale, int tp_rank, int blocksparse_local_blocks,
                                   int blocksparse_vert_stride, int blocksparse_block_size,
                                   int blocksparse_head_sliding_step, void* stream);

/* ---------------------------------------------------------------------------------------------
 * gptq_marlin_gemm - replaces kernels/quantization/gptq_marlin/gptq_marlin.cu:2263-2468.
 *   c            [size_m, size_n] scalar dtype (written; caller-allocated here, the torch
 *                binding allocates and returns it like the reference :2336-2341)
 *   a            [size_m, size_k] scalar dtype, row stride lda elements (lda % 8 == 0)
 *   b_q_weight   [size_k/16, size_n*16/pack] int32, Marlin layout (marlin_utils_test.py:30-92)
 *   b_scales     [num_groups, size_n] scalar dtype, Marlin-permuted (marlin_utils.py:204-225)
 *   b_zeros      [num_groups, size_n/pack] int32 or NULL (has_zp)
 *   g_idx, perm  [size_k] int32 or NULL (act-order)
 *   workspace    the reference's int32 lock buffer (>= size_n/64*16 ints); left all-zero
 *   b_q_type     ScalarType id (kernels/core/scalar_type.hpp:100-138): uint4b8, uint8b128,
 *                uint4 (has_zp), uint8 (has_zp)
 *   scratch      >= aph_gptq_marlin_gemm_scratch_bytes(...) bytes (fp32 split-K partials)
 * ------------------------------------------------------------------------------------------- */
size_t aph_gptq_marlin_gemm_scratch_bytes(int64_t size_m, int64_t size_n, int64_t size_k);

int aph_gptq_marlin_gemm(void* c, const void* a, const void* b_q_weight, const void* b_scales,
                         const void* b_zeros, const int32_t* g_idx, const int32_t* perm, int32_t* workspace,
                         int64_t workspace_ints, int64_t b_q_type, int64_t size_m, int64_t size_n,
                         int64_t size_k, int64_t lda, int64_t num_groups, int is_k_full, int has_zp,
                         int use_atomic_add, int use_fp32_reduce, int is_zp_float, int dtype, void* scratch,
                         size_t scratch_bytes, void* stream);

/* MLA cache ops (rest of kernels/cache_kernels.cu).
 *   aph_copy_blocks_mla       cache_kernels.cu:100-117,173-208: aph_copy_blocks over ONE joint cache per layer;
 *                             block_bytes = stride(0) * element size (padding included)
 *   aph_concat_and_cache_mla  cache_kernels.cu:308-348,463-500: kv_cache[slot] = [kv_c[t] | k_pe[t]] (slot < 0: skipped),
 *                             FP8-encoded with *scale unless kv_dtype is APH_KV_AUTO; strides in elements
 *   aph_gather_cache          cache_kernels.cu:582-735: dst[cu_seq_lens[b] + t] = cache[block_table[b][seq_starts[b] /
 *                             block_size + t / block_size]][t % block_size]; seq_starts may be NULL; strides in BYTES
 */
int aph_copy_blocks_mla(const int64_t* kv_cache_ptrs, const int64_t* block_mapping, int num_layers, int64_t num_pairs,
                        int64_t block_bytes, void* stream);
int aph_concat_and_cache_mla(const void* kv_c, const void* k_pe, void* kv_cache, const int64_t* slot_mapping,
                             int64_t num_tokens, int kv_lora_rank, int pe_dim, int block_size, int64_t block_stride,
                             int64_t entry_stride, int64_t kv_c_stride, int64_t k_pe_stride, int dtype, int kv_dtype,
                             const float* scale, void* stream);
int aph_gather_cache(const void* src_cache, void* dst, const int32_t* block_table, const int32_t* cu_seq_lens,
                     const int32_t* seq_starts, int batch_size, int block_size, int64_t entry_bytes,
                     int64_t block_table_stride, int64_t cache_block_stride_bytes, int64_t cache_entry_stride_bytes,
                     int64_t dst_entry_stride_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Marlin repack - replaces kernels/quantization/gptq_marlin/gptq_marlin_repack.cu:236-322 and
 * awq_marlin_repack.cu:189-268: checkpoint formats -> the Marlin wire format gptq_marlin_gemm reads.
 *   out         [size_k / 16, size_n * 16 / pack] int32, pack = 32 / num_bits (4 or 8 bits)
 *   GPTQ input  [size_k / pack, size_n] int32 (packed along k); perm: int32 [size_k] act-order row order
 *               (output row k reads input row perm[k]) or NULL
 *   AWQ input   [size_k, size_n / pack] int32 (packed along n with AWQ's column interleave)
 * size_k % 16 == 0 and size_n % 64 == 0. Bit-identical with the reference kernels.
 * ------------------------------------------------------------------------------------------- */
/* Act-order (desc_act) support - replaces permute_cols_kernel, gptq_marlin.cu:459-515,2187-2194: out[m][k] =
 * a[m][perm[k]] for 2-byte element types, out dense [size_m, size_k]. With weights repacked with the same `perm`
 * (rows sorted by group) and is_k_full, act-order GEMM = aph_permute_cols + aph_gptq_marlin_gemm with g_idx = perm =
 * NULL; that is what the torch op does. */
int aph_permute_cols(void* out, const void* a, const int32_t* perm, int64_t size_m, int64_t size_k, int64_t lda,
                     void* stream);
int aph_gptq_marlin_repack(int32_t* out, const int32_t* b_q_weight, const int32_t* perm, int64_t size_k,
                           int64_t size_n, int num_bits, void* stream);
int aph_awq_marlin_repack(int32_t* out, const int32_t* b_q_weight, int64_t size_k, int64_t size_n, int num_bits,
                          void* stream);

/* ---------------------------------------------------------------------------------------------
 * Decode-layer glue - replaces kernels/layernorm_kernels.cu:139-240, activation_kernels.cu:82-86
 * and pos_encoding_kernels.cu:124-196. dtype: aph_dtype_t of every tensor of the call. Row-major, last dimension
 * contiguous. The arithmetic rounds to the storage type exactly where the reference does.
 *   aph_rms_norm            out[t] = T(x[t] * rsqrt(mean(x[t]^2) + eps)) * weight;  input rows input_stride apart
 *   aph_fused_add_rms_norm  residual[t] += input[t];  input[t] = norm(residual[t]) * weight   (both in place)
 *   aph_silu_and_mul        out[t][i] = silu(input[t][i]) * input[t][d + i],  input [tokens, 2 d], at most 65535 tokens
 *   aph_rotary_embedding    rotates the first rot_dim elements of every head of query (and key unless NULL) in place
 *                           by cos_sin_cache[positions[t]] ([max_position, rot_dim]: cos half, sin half);
 *                           is_neox: pairs (i, i + rot_dim/2), else (2i, 2i+1); strides in elements per token
 * ------------------------------------------------------------------------------------------- */
int aph_rms_norm(void* out, const void* input, const void* weight, float epsilon, int64_t num_tokens,
                 int64_t hidden_size, int64_t input_stride, int dtype, void* stream);
int aph_fused_add_rms_norm(void* input, void* residual, const void* weight, float epsilon, int64_t num_tokens,
                           int64_t hidden_size, int dtype, void* stream);
int aph_silu_and_mul(void* out, const void* input, int64_t num_tokens, int64_t d, int dtype, void* stream);
/* The other activations of kernels/activation_kernels.cu:82-225. Gated (input [tokens, 2 d] -> out [tokens, d]):
 * silu_and_mul, mul_and_silu (activation on the second half), gelu_and_mul (erf), gelu_tanh_and_mul, fatrelu_and_mul
 * (param = threshold). Element-wise over numel values: gelu_new, gelu_fast, gelu_quick. */
typedef enum {
  APH_ACT_SILU_AND_MUL = 0, APH_ACT_MUL_AND_SILU = 1, APH_ACT_GELU_AND_MUL = 2, APH_ACT_GELU_TANH_AND_MUL = 3,
  APH_ACT_FATRELU_AND_MUL = 4, APH_ACT_GELU_NEW = 5, APH_ACT_GELU_FAST = 6, APH_ACT_GELU_QUICK = 7
} aph_activation_t;
int aph_act_and_mul(int act, void* out, const void* input, int64_t num_tokens, int64_t d, float param, int dtype,
                    void* stream);
int aph_activation(int act, void* out, const void* input, int64_t numel, int dtype, void* stream);
/* cos_sin_cache_offsets: NULL, or int64 [num_tokens] added to positions (batched_rotary_embedding,
 * pos_encoding_kernels.cu:98-121,204-260) */
int aph_rotary_embedding(const int64_t* positions, void* query, void* key, const void* cos_sin_cache,
                         const int64_t* cos_sin_cache_offsets, int64_t num_tokens, int rot_dim, int64_t query_stride, int64_t key_stride, int num_heads,
                         int num_kv_heads, int head_size, int is_neox, int dtype, void* stream);

/* Multi-step decode bookkeeping - replaces kernels/prepare_inputs/advance_step.cu:12-66,190-232
 * (advance_step_flashattn): for q < num_queries: input_tokens[q] = sampled_token_ids[q]; seq_lens[q] += 1;
 * input_positions[q] = seq_lens[q] - 1; slot_mapping[q] = block_tables[q][pos / block_size] * block_size + pos % block_size;
 * rows num_queries..num_seqs-1 (CUDA-graph padding) get token 0, position 0, slot -1. */
int aph_advance_step_flashattn(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                               const int64_t* sampled_token_ids, int64_t* input_positions, int32_t* seq_lens,
                               int64_t* slot_mapping, const int32_t* block_tables, int64_t block_tables_stride,
                               void* stream);

/* The FlashInfer flavour - replaces kernels/prepare_inputs/advance_step.cu:91-167,234-311 (advance_step_flashinfer):
 * the updates above, plus paged_kv_last_page_len[q] = pos % block_size + 1, block_table_bound[q] =
 * ceil(seq_lens[q] / block_size), paged_kv_indptr[0] = 0, paged_kv_indptr[q + 1] = sum of block_table_bound[0..q],
 * paged_kv_indptr[num_queries + 1 .. num_seqs] = paged_kv_indptr[num_queries], and paged_kv_indices[indptr[q] + c] =
 * block_tables[q][c] for c < block_table_bound[q]. paged_kv_indices_len is the capacity of paged_kv_indices in
 * elements: writes beyond it are dropped (the reference writes out of bounds). Two launches on `stream`. */
int aph_advance_step_flashinfer(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                                const int64_t* sampled_token_ids, int64_t* input_positions, int32_t* seq_lens,
                                int64_t* slot_mapping, const int32_t* block_tables, int64_t block_tables_stride,
                                int32_t* paged_kv_indices, int64_t paged_kv_indices_len, int32_t* paged_kv_indptr,
                                int32_t* paged_kv_last_page_len, int32_t* block_table_bound, void* stream);

/* ---------------------------------------------------------------------------------------------
 * One-shot all-reduce over NVLink peer memory - replaces cross_device_reduce_1stage,
 * kernels/all_reduce/custom_all_reduce.cuh:296-318 (barriers :140-200, dispatch :525-600).
 * Every rank owns one symmetric buffer of aph_all_reduce_buffer_bytes(world, slot_bytes) bytes laid out
 * [aph_all_reduce_signal_bytes() of flags, zeroed once | 2 halves x world slots x slot_bytes], mapped into every peer
 * (CUDA IPC / torch symmetric memory); peer_bufs_dev is a DEVICE array of `world` base pointers in rank order; counters
 * is a rank-local zeroed uint32[aph_all_reduce_max_blocks()]. Each rank pushes its input into its slot on every rank,
 * one flag barrier, then out[i] = sum over slots, added in fp32 in rank order and rounded once (bit-identical on every
 * rank); out may alias in. numel * element size must be a multiple of 16 and at most slot_bytes. All ranks must call
 * with the same numel, dtype and max_blocks, in the same order. world in {2, 4, 6, 8}.
 * ------------------------------------------------------------------------------------------- */
size_t aph_all_reduce_signal_bytes(void);
size_t aph_all_reduce_buffer_bytes(int world, int64_t slot_bytes);
int aph_all_reduce_max_blocks(void);
int aph_one_shot_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank,
                            int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream);

/* Two-shot variant - replaces cross_device_reduce_2stage, custom_all_reduce.cuh:321-400 (dispatch :563-570). Same buffers,
 * counters, arguments and result (bit-identical on every rank) as aph_one_shot_all_reduce, and the two may be mixed
 * freely call by call. Every 16-byte vector has an owner rank: partials go to the owner only, the owner adds them in
 * rank order in fp32, rounds once and hands the sum to every rank. NVLink egress per rank: 2 (world - 1) / world x the
 * message instead of (world - 1) x, for one more flag barrier: the choice for larger groups and messages. */
int aph_two_shot_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank,
                            int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream);

/* aph_one_shot_all_reduce fused with the residual add + RMSNorm that follows every all-reduce of a decoder layer
 * (aph_fused_add_rms_norm): x [num_tokens, hidden] holds this rank's partial on entry and the normalised sum on exit,
 * residual += sum (rounded as the unfused pair rounds). Same buffers, counters and calling discipline as
 * aph_one_shot_all_reduce (the two may be mixed freely); float16 / bfloat16, hidden_size % 8 == 0 and <= 16384. */
int aph_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon,
                                const void* const* peer_bufs_dev, uint32_t* counters, int rank, int world,
                                int64_t num_tokens, int64_t hidden_size, int dtype, int64_t slot_bytes, void* stream);
/* the same on top of aph_two_shot_all_reduce (identical results) */
int aph_two_shot_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon,
                                         const void* const* peer_bufs_dev, uint32_t* counters, int rank, int world,
                                         int64_t num_tokens, int64_t hidden_size, int dtype, int64_t slot_bytes, void* stream);

/* The LL variants: the same two-shot exchange with NCCL's low-latency wire format - every 8 bytes carry 4 bytes of data
 * and the call number as a flag in one atomic store, so the arrival of the data is the signal and there is no flag
 * barrier and no fence (two one-way NVLink latencies instead of two barriers; twice the bytes on the wire). Messages up
 * to aph_all_reduce_ll_max_bytes(); the fused form also needs num_tokens * ceil(hidden_size / 4096) <=
 * aph_all_reduce_ll_max_passes(). Rank-order fp32 sums: bit-identical to the one-shot / two-shot calls, with which they
 * mix freely. The buffer (aph_all_reduce_buffer_bytes) must be ZEROED once before the first call. */
size_t aph_all_reduce_ll_max_bytes(void);
int aph_all_reduce_ll_max_passes(void);
int aph_ll_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank,
                      int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream);
int aph_ll_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon,
                                   const void* const* peer_bufs_dev, uint32_t* counters, int rank, int world,
                                   int64_t num_tokens, int64_t hidden_size, int dtype, int64_t slot_bytes, void* stream);

/* The NVLS ("multimem") variants - the reference has no counterpart (its custom all-reduce predates NVSwitch reductions;
 * NCCL's NVLS algorithm is what it would fall back to): multicast_buf is the MULTICAST mapping of the same symmetric
 * buffer (one address = the same offset in every rank's buffer; torch symmetric memory: handle.multicast_ptr). Each
 * 16-byte vector is summed inside the switch by one multimem.ld_reduce of its owner rank and handed to every rank by one
 * multimem.st. Same buffers, counters and calling discipline as the calls above (all may be mixed freely). The result is
 * identical on every rank and from run to run; the order of the additions is the switch's, so it may differ from the
 * rank-order sum of the other two variants in the last bit. Needs a device and driver with multicast support. */
int aph_multimem_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, void* multicast_buf,
                            uint32_t* counters, int rank, int world, int64_t numel, int dtype, int64_t slot_bytes,
                            int max_blocks, void* stream);
int aph_multimem_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon,
                                         const void* const* peer_bufs_dev, void* multicast_buf, uint32_t* counters,
                                         int rank, int world, int64_t num_tokens, int64_t hidden_size, int dtype,
                                         int64_t slot_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* APHRODITE_B200_H_ */
