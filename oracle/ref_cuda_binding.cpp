// TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.
//
// Trimmed binding that exposes the UNMODIFIED reference CUDA kernels of the hot path
//   /root/reference/kernels/attention/paged_attention_v1.cu  (:170-192)
//   /root/reference/kernels/attention/paged_attention_v2.cu  (:176-201)
//   /root/reference/kernels/cache_kernels.cu                 (reshape_and_cache :369-395 and the
//                                                             other cache ops of that file)
//   /root/reference/kernels/quantization/gptq_marlin/gptq_marlin.cu        (:2263-2474)
//   /root/reference/kernels/quantization/gptq_marlin/gptq_marlin_repack.cu
//   /root/reference/kernels/quantization/gptq_marlin/awq_marlin_repack.cu
// under the private dispatcher namespace `_refcuda`, with the schema strings of
// kernels/torch_bindings.cpp:41-65,326-347,680-734. The Marlin translation units register their
// own CUDA impls into TORCH_EXTENSION_NAME (= _refcuda, set by oracle/Makefile). The sources are
// compiled where they lie (oracle/Makefile, target `refcuda`, nvcc for sm_100a); nothing is copied
// into this repository. It is the same-hardware parity target and "kernel to beat" on the GPU box.
#include <torch/all.h>
#include <torch/library.h>
#include <optional>
#include <string>
#include <vector>

#define ATTN_TAIL                                                                                  \
  torch::Tensor &query, torch::Tensor &key_cache, torch::Tensor &value_cache, int64_t num_kv_heads, \
      double scale, torch::Tensor &block_tables, torch::Tensor &seq_lens, int64_t block_size,      \
      int64_t max_seq_len, const std::optional<torch::Tensor> &alibi_slopes,                       \
      const std::string &kv_cache_dtype, torch::Tensor &k_scale, torch::Tensor &v_scale,           \
      const int64_t tp_rank, const int64_t blocksparse_local_blocks,                               \
      const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,                 \
      const int64_t blocksparse_head_sliding_step

void paged_attention_v1(torch::Tensor& out, ATTN_TAIL);
void paged_attention_v2(torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
                        torch::Tensor& tmp_out, ATTN_TAIL);
void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const torch::Tensor& block_mapping);
void copy_blocks(std::vector<torch::Tensor> const& key_caches, std::vector<torch::Tensor> const& value_caches,
                 const torch::Tensor& block_mapping);
void copy_blocks_mla(std::vector<torch::Tensor> const& kv_caches, const torch::Tensor& block_mapping);
void advance_step_flashattn(int64_t num_seqs, int64_t num_queries, int64_t block_size, torch::Tensor& input_tokens,
                            torch::Tensor& sampled_token_ids, torch::Tensor& input_positions, torch::Tensor& seq_lens,
                            torch::Tensor& slot_mapping, torch::Tensor& block_tables);
void advance_step_flashinfer(int64_t num_seqs, int64_t num_queries, int64_t block_size, torch::Tensor& input_tokens,
                             torch::Tensor& sampled_token_ids, torch::Tensor& input_positions, torch::Tensor& seq_lens,
                             torch::Tensor& slot_mapping, torch::Tensor& block_tables, torch::Tensor& paged_kv_indices,
                             torch::Tensor& paged_kv_indptr, torch::Tensor& paged_kv_last_page_len,
                             torch::Tensor& block_table_bounds);
void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                       torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype, torch::Tensor& k_scale, torch::Tensor& v_scale);
void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                             torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                             const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                             torch::Tensor& v_scale);
void concat_and_cache_mla(torch::Tensor& kv_c, torch::Tensor& k_pe, torch::Tensor& kv_cache,
                          torch::Tensor& slot_mapping, const std::string& kv_cache_dtype,
                          torch::Tensor& scale);
void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache, const double scale,
                 const std::string& kv_cache_dtype);
void gather_cache(torch::Tensor const& src_cache, torch::Tensor const& dst, torch::Tensor const& block_table,
                  torch::Tensor const& cu_seq_lens, int64_t batch_size, std::optional<torch::Tensor> seq_starts);

// kernels/layernorm_kernels.cu:139,168; activation_kernels.cu:82; pos_encoding_kernels.cu:124
void rms_norm(torch::Tensor& out, torch::Tensor& input, torch::Tensor& weight, double epsilon);
void fused_add_rms_norm(torch::Tensor& input, torch::Tensor& residual, torch::Tensor& weight, double epsilon);
void silu_and_mul(torch::Tensor& out, torch::Tensor& input);
void mul_and_silu(torch::Tensor& out, torch::Tensor& input);
void gelu_and_mul(torch::Tensor& out, torch::Tensor& input);
void gelu_tanh_and_mul(torch::Tensor& out, torch::Tensor& input);
void fatrelu_and_mul(torch::Tensor& out, torch::Tensor& input, double threshold);
void gelu_new(torch::Tensor& out, torch::Tensor& input);
void gelu_fast(torch::Tensor& out, torch::Tensor& input);
void gelu_quick(torch::Tensor& out, torch::Tensor& input);
void batched_rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int64_t head_size,
                              torch::Tensor& cos_sin_cache, bool is_neox, int64_t rot_dim,
                              torch::Tensor& cos_sin_cache_offsets);
void rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int64_t head_size,
                      torch::Tensor& cos_sin_cache, bool is_neox);

#define ATTN_SCHEMA_TAIL                                                                            \
  "Tensor query, Tensor key_cache, Tensor value_cache, int num_kv_heads, float scale, "            \
  "Tensor block_tables, Tensor seq_lens, int block_size, int max_seq_len, Tensor? alibi_slopes, "   \
  "str kv_cache_dtype, Tensor k_scale, Tensor v_scale, int tp_rank, int blocksparse_local_blocks, " \
  "int blocksparse_vert_stride, int blocksparse_block_size, int blocksparse_head_sliding_step) -> ()"

TORCH_LIBRARY(_refcuda, ops) {
  ops.def("paged_attention_v1(Tensor! out, " ATTN_SCHEMA_TAIL);
  ops.impl("paged_attention_v1", torch::kCUDA, &paged_attention_v1);
  ops.def("paged_attention_v2(Tensor! out, Tensor! exp_sums, Tensor! max_logits, Tensor! tmp_out, " ATTN_SCHEMA_TAIL);
  ops.impl("paged_attention_v2", torch::kCUDA, &paged_attention_v2);
  // impls registered by the Marlin translation units themselves
  ops.def(
      "gptq_marlin_gemm(Tensor a, Tensor b_q_weight, Tensor b_scales, Tensor b_zeros, Tensor g_idx, Tensor perm, "
      "Tensor workspace, int b_q_type, SymInt size_m, SymInt size_n, SymInt size_k, bool is_k_full, bool has_zp, "
      "bool use_atomic_add, bool use_fp32_reduce, bool is_zp_float) -> Tensor");
  ops.def("gptq_marlin_repack(Tensor b_q_weight, Tensor perm, SymInt size_k, SymInt size_n, int num_bits) -> Tensor");
  ops.def("awq_marlin_repack(Tensor b_q_weight, SymInt size_k, SymInt size_n, int num_bits) -> Tensor");
  // cache ops (kernels/torch_bindings.cpp:680-734)
  ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  ops.impl("swap_blocks", torch::kCUDA, &swap_blocks);
  ops.def("copy_blocks(Tensor(a!)[] key_caches, Tensor[](b!) value_caches, Tensor block_mapping) -> ()");
  ops.impl("copy_blocks", torch::kCUDA, &copy_blocks);
  ops.def(
      "advance_step_flashattn(int num_seqs, int num_queries, int block_size, Tensor! input_tokens, Tensor sampled_token_ids, "
      "Tensor! input_positions, Tensor! seq_lens, Tensor! slot_mapping, Tensor block_tables) -> ()");
  ops.impl("advance_step_flashattn", torch::kCUDA, &advance_step_flashattn);
  ops.def(
      "advance_step_flashinfer(int num_seqs, int num_queries, int block_size, Tensor! input_tokens, Tensor sampled_token_ids, "
      "Tensor! input_positions, Tensor! seq_lens, Tensor! slot_mapping, Tensor block_tables, Tensor! paged_kv_indices, "
      "Tensor! paged_kv_indptr, Tensor! paged_kv_last_page_len, Tensor! block_table_bounds) -> ()");
  ops.impl("advance_step_flashinfer", torch::kCUDA, &advance_step_flashinfer);
  ops.def("copy_blocks_mla(Tensor(a!)[] kv_caches, Tensor block_mapping) -> ()");
  ops.impl("copy_blocks_mla", torch::kCUDA, &copy_blocks_mla);
  ops.def(
      "reshape_and_cache(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, Tensor slot_mapping, "
      "str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()");
  ops.impl("reshape_and_cache", torch::kCUDA, &reshape_and_cache);
  ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value, Tensor! key_cache, Tensor! value_cache, "
      "Tensor slot_mapping, str kv_cache_dtype, Tensor k_scale, Tensor v_scale) -> ()");
  ops.impl("reshape_and_cache_flash", torch::kCUDA, &reshape_and_cache_flash);
  ops.def(
      "concat_and_cache_mla(Tensor kv_c, Tensor k_pe, Tensor! kv_cache, Tensor slot_mapping, str kv_cache_dtype, "
      "Tensor scale) -> ()");
  ops.impl("concat_and_cache_mla", torch::kCUDA, &concat_and_cache_mla);
  ops.def("convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, str kv_cache_dtype) -> ()");
  ops.impl("convert_fp8", torch::kCUDA, &convert_fp8);
  ops.def(
      "gather_cache(Tensor src_cache, Tensor! dst, Tensor block_table, Tensor cu_seq_lens, int batch_size, "
      "Tensor? seq_starts) -> ()");
  ops.impl("gather_cache", torch::kCUDA, &gather_cache);
  ops.def("rms_norm(Tensor! result, Tensor input, Tensor weight, float epsilon) -> ()");
  ops.impl("rms_norm", torch::kCUDA, &rms_norm);
  ops.def("fused_add_rms_norm(Tensor! input, Tensor! residual, Tensor weight, float epsilon) -> ()");
  ops.impl("fused_add_rms_norm", torch::kCUDA, &fused_add_rms_norm);
  ops.def("silu_and_mul(Tensor! result, Tensor input) -> ()");
  ops.impl("silu_and_mul", torch::kCUDA, &silu_and_mul);
  ops.def("mul_and_silu(Tensor! out, Tensor input) -> ()");
  ops.impl("mul_and_silu", torch::kCUDA, &mul_and_silu);
  ops.def("gelu_and_mul(Tensor! out, Tensor input) -> ()");
  ops.impl("gelu_and_mul", torch::kCUDA, &gelu_and_mul);
  ops.def("gelu_tanh_and_mul(Tensor! out, Tensor input) -> ()");
  ops.impl("gelu_tanh_and_mul", torch::kCUDA, &gelu_tanh_and_mul);
  ops.def("fatrelu_and_mul(Tensor! out, Tensor input, float threshold) -> ()");
  ops.impl("fatrelu_and_mul", torch::kCUDA, &fatrelu_and_mul);
  ops.def("gelu_new(Tensor! out, Tensor input) -> ()");
  ops.impl("gelu_new", torch::kCUDA, &gelu_new);
  ops.def("gelu_fast(Tensor! out, Tensor input) -> ()");
  ops.impl("gelu_fast", torch::kCUDA, &gelu_fast);
  ops.def("gelu_quick(Tensor! out, Tensor input) -> ()");
  ops.impl("gelu_quick", torch::kCUDA, &gelu_quick);
  ops.def(
      "batched_rotary_embedding(Tensor positions, Tensor! query, Tensor! key, int head_size, Tensor cos_sin_cache, "
      "bool is_neox, int rot_dim, Tensor cos_sin_cache_offsets) -> ()");
  ops.impl("batched_rotary_embedding", torch::kCUDA, &batched_rotary_embedding);
  ops.def(
      "rotary_embedding(Tensor positions, Tensor! query, Tensor! key, int head_size, Tensor cos_sin_cache, "
      "bool is_neox) -> ()");
  ops.impl("rotary_embedding", torch::kCUDA, &rotary_embedding);
}
