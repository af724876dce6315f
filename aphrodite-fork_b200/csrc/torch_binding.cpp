// Thin torch C++ extension over the C ABI (include/aphrodite_b200.h).
//
// Registers CUDA implementations for exactly the reference's four op schemas
// (/root/reference/kernels/torch_bindings.cpp:41-51, :54-65, :326-333, :696-702) with the
// reference's C++ signatures (kernels/ops.h:32-54, kernels/cache.h:21-25,
// kernels/quantization/gptq_marlin/gptq_marlin.cu:2263-2270).
//
// Two build flavours of this one file (SURVEY.md 8b "Loading into a live Aphrodite"):
//   -DAPH_DEFINE_SCHEMAS  standalone: also defines the schemas (TORCH_LIBRARY_FRAGMENT), for
//                         processes where aphrodite._C is not loaded (the GPU test box);
//   (no flag)             overlay: impl-only registration on the BackendSelect key of an already-imported
//                         aphrodite._C: CUDA calls run here, argument combinations this library does not serve
//                         fall through to the reference's own CUDA kernel (bottom of this file).
// convert_fp8 (cache_kernels.cu:499-577) is registered too because the reference's tests and
// create_kv_caches_with_random depend on it.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/all.h>
#include <torch/library.h>
#include <ATen/core/dispatch/Dispatcher.h>

#include <optional>
#include <string>
#include <vector>

#include "../../include/aphrodite_b200.h"

namespace {

// A combination of arguments the reference serves and this library does not (APH_ERR_UNSUPPORTED; nothing has been
// launched when it is reported). Standalone it surfaces as the RuntimeError it is; in the overlay flavour the op's
// wrapper catches it and hands the call to the reference's own kernel (see Overlay below).
struct UnsupportedHere : public c10::Error {
  using c10::Error::Error;
};

inline void check_status(int st, const char* op) {
  if (st == APH_ERR_UNSUPPORTED)
    throw UnsupportedHere({__func__, __FILE__, static_cast<uint32_t>(__LINE__)}, std::string(op) + ": " + aph_b200_last_error());
  TORCH_CHECK(st == APH_OK, op, ": ", aph_b200_last_error());
}

inline int dtype_code(const torch::Tensor& t, const char* op) {
  switch (t.scalar_type()) {
    case at::ScalarType::Float: return APH_DT_F32;
    case at::ScalarType::Half: return APH_DT_F16;
    case at::ScalarType::BFloat16: return APH_DT_BF16;
    default: TORCH_CHECK(false, op, ": unsupported data type ", t.scalar_type());
  }
  return -1;
}

inline int kv_code(const std::string& s) {
  const int kv = aph_kv_dtype_from_string(s.c_str());
  TORCH_CHECK(kv >= 0, "Unsupported data type of kv cache: ", s);
  return kv;
}

inline void* stream_of() { return (void*)at::cuda::getCurrentCUDAStream().stream(); }

void paged_attention_impl(bool v2, torch::Tensor& out, torch::Tensor* exp_sums, torch::Tensor* max_logits,
                          torch::Tensor* tmp_out, torch::Tensor& query, torch::Tensor& key_cache,
                          torch::Tensor& value_cache, int64_t num_kv_heads, double scale, torch::Tensor& block_tables,
                          torch::Tensor& seq_lens, int64_t block_size, int64_t max_seq_len,
                          const std::optional<torch::Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                          torch::Tensor& k_scale, torch::Tensor& v_scale, int64_t tp_rank,
                          int64_t blocksparse_local_blocks, int64_t blocksparse_vert_stride,
                          int64_t blocksparse_block_size, int64_t blocksparse_head_sliding_step) {
  const char* op = v2 ? "paged_attention_v2" : "paged_attention_v1";
  TORCH_CHECK(query.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda() && out.is_cuda(), op,
              ": tensors must be on a CUDA device");
  TORCH_CHECK(query.dim() == 3 && out.dim() == 3, op, ": query/out must be [num_seqs, num_heads, head_size]");
  TORCH_CHECK(query.stride(2) == 1 && query.stride(1) == query.size(2), op, ": query heads must be contiguous");
  TORCH_CHECK(out.stride(2) == 1 && out.stride(1) == out.size(2), op, ": out heads must be contiguous");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && seq_lens.scalar_type() == at::ScalarType::Int, op,
              ": block_tables and seq_lens must be int32");
  TORCH_CHECK(block_tables.dim() == 2 && block_tables.stride(1) == 1, op, ": block_tables must be 2-D, row-major");
  TORCH_CHECK(k_scale.scalar_type() == at::ScalarType::Float && v_scale.scalar_type() == at::ScalarType::Float, op,
              ": k_scale / v_scale must be float32 tensors");
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const int kv = kv_code(kv_cache_dtype);
  const int dt = dtype_code(query, op);
  const int num_seqs = (int)query.size(0), num_heads = (int)query.size(1), head_size = (int)query.size(2);
  const float* alibi = nullptr;
  if (alibi_slopes.has_value() && alibi_slopes->defined()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::ScalarType::Float && alibi_slopes->is_contiguous(), op,
                ": alibi_slopes must be contiguous float32");
    alibi = alibi_slopes->data_ptr<float>();
  }
  const size_t ws_bytes =
      aph_paged_attention_workspace_bytes(num_seqs, num_heads, (int)num_kv_heads, head_size, (int)max_seq_len);
  torch::Tensor ws;
  void* ws_ptr = nullptr;
  if (ws_bytes > 0) {
    // torch caching allocator: stream-ordered and CUDA-graph safe, like the reference's temporaries
    ws = torch::empty({(int64_t)ws_bytes}, torch::TensorOptions().dtype(torch::kUInt8).device(query.device()));
    ws_ptr = ws.data_ptr();
  }
  int st;
  if (!v2) {
    st = aph_paged_attention_v1(out.data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                num_seqs, num_heads, (int)num_kv_heads, head_size, (float)scale,
                                block_tables.data_ptr<int32_t>(), seq_lens.data_ptr<int32_t>(), (int)block_size,
                                (int)max_seq_len, (int)block_tables.size(1), alibi, query.stride(0), out.stride(0),
                                key_cache.stride(0), key_cache.stride(1), dt, kv, k_scale.data_ptr<float>(),
                                v_scale.data_ptr<float>(), (int)tp_rank, (int)blocksparse_local_blocks,
                                (int)blocksparse_vert_stride, (int)blocksparse_block_size,
                                (int)blocksparse_head_sliding_step, ws_ptr, ws_bytes, stream_of());
  } else {
    st = aph_paged_attention_v2(out.data_ptr(), exp_sums->data_ptr<float>(), max_logits->data_ptr<float>(),
                                tmp_out->data_ptr(), query.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(),
                                num_seqs, num_heads, (int)num_kv_heads, head_size, (float)scale,
                                block_tables.data_ptr<int32_t>(), seq_lens.data_ptr<int32_t>(), (int)block_size,
                                (int)max_seq_len, (int)block_tables.size(1), alibi, query.stride(0), out.stride(0),
                                key_cache.stride(0), key_cache.stride(1), dt, kv, k_scale.data_ptr<float>(),
                                v_scale.data_ptr<float>(), (int)tp_rank, (int)blocksparse_local_blocks,
                                (int)blocksparse_vert_stride, (int)blocksparse_block_size,
                                (int)blocksparse_head_sliding_step, ws_ptr, ws_bytes, stream_of());
  }
  check_status(st, op);
}

void paged_attention_v1(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache, torch::Tensor& value_cache,
                        int64_t num_kv_heads, double scale, torch::Tensor& block_tables, torch::Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, torch::Tensor& k_scale, torch::Tensor& v_scale,
                        const int64_t tp_rank, const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  paged_attention_impl(false, out, nullptr, nullptr, nullptr, query, key_cache, value_cache, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale,
                       tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                       blocksparse_head_sliding_step);
}

void paged_attention_v2(torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits, torch::Tensor& tmp_out,
                        torch::Tensor& query, torch::Tensor& key_cache, torch::Tensor& value_cache,
                        int64_t num_kv_heads, double scale, torch::Tensor& block_tables, torch::Tensor& seq_lens,
                        int64_t block_size, int64_t max_seq_len, const std::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype, torch::Tensor& k_scale, torch::Tensor& v_scale,
                        const int64_t tp_rank, const int64_t blocksparse_local_blocks,
                        const int64_t blocksparse_vert_stride, const int64_t blocksparse_block_size,
                        const int64_t blocksparse_head_sliding_step) {
  paged_attention_impl(true, out, &exp_sums, &max_logits, &tmp_out, query, key_cache, value_cache, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale,
                       tp_rank, blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                       blocksparse_head_sliding_step);
}

void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache, torch::Tensor& value_cache,
                       torch::Tensor& slot_mapping, const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                       torch::Tensor& v_scale) {
  const char* op = "reshape_and_cache";
  TORCH_CHECK(key.is_cuda() && value.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda(), op,
              ": tensors must be on a CUDA device");
  TORCH_CHECK(key.dim() == 3 && value.dim() == 3 && key_cache.dim() == 5 && value_cache.dim() == 4, op,
              ": bad tensor ranks");
  TORCH_CHECK(key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 &&
                  value.stride(1) == value.size(2),
              op, ": key/value heads must be contiguous");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(), op,
              ": slot_mapping must be contiguous int64");
  TORCH_CHECK(k_scale.scalar_type() == at::ScalarType::Float && v_scale.scalar_type() == at::ScalarType::Float, op,
              ": k_scale / v_scale must be float32 tensors");
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  const int kv = kv_code(kv_cache_dtype);
  // reference: num_tokens = slot_mapping.size(0) (cache_kernels.cu:379)
  const int st = aph_reshape_and_cache(
      key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr<int64_t>(),
      slot_mapping.size(0), key.stride(0), value.stride(0), (int)key.size(1), (int)key.size(2),
      (int)key_cache.size(3), (int)key_cache.size(4), key_cache.stride(0), dtype_code(key, op), kv,
      k_scale.data_ptr<float>(), v_scale.data_ptr<float>(), stream_of());
  check_status(st, op);
}

void convert_fp8(torch::Tensor& dst_cache, torch::Tensor& src_cache, const double scale,
                 const std::string& kv_cache_dtype) {
  const char* op = "convert_fp8";
  TORCH_CHECK(src_cache.is_cuda() && dst_cache.is_cuda() && src_cache.device() == dst_cache.device(), op,
              ": src and dst must be on the same CUDA device");
  TORCH_CHECK(src_cache.is_contiguous() && dst_cache.is_contiguous() && src_cache.numel() == dst_cache.numel(), op,
              ": src and dst must be contiguous and of equal size");
  const at::cuda::OptionalCUDAGuard guard(device_of(src_cache));
  int kv = kv_code(kv_cache_dtype);
  if (kv == APH_KV_AUTO) kv = APH_KV_FP8_E4M3;  // reference: "auto" behaves as e4m3 here (cache_kernels.cu:540-556)
  int st;
  if (dst_cache.scalar_type() == at::ScalarType::Byte) {
    st = aph_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(), (float)scale,
                         dtype_code(src_cache, op), kv, 0, stream_of());
  } else {
    TORCH_CHECK(src_cache.scalar_type() == at::ScalarType::Byte, op, ": one of src/dst must be uint8");
    st = aph_convert_fp8(dst_cache.data_ptr(), src_cache.data_ptr(), src_cache.numel(), (float)scale,
                         dtype_code(dst_cache, op), kv, 1, stream_of());
  }
  check_status(st, op);
}

// argument checks of the reference entry point, gptq_marlin.cu:2271-2332
// ---- the rest of cache_kernels.cu a block manager calls ----
void reshape_and_cache_flash(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache, torch::Tensor& value_cache,
                             torch::Tensor& slot_mapping, const std::string& kv_cache_dtype, torch::Tensor& k_scale,
                             torch::Tensor& v_scale) {
  const char* op = "reshape_and_cache_flash";
  TORCH_CHECK(key.is_cuda() && value.is_cuda() && key_cache.is_cuda() && value_cache.is_cuda(), op,
              ": tensors must be on a CUDA device");
  TORCH_CHECK(key.dim() == 3 && value.dim() == 3 && key_cache.dim() == 4 && value_cache.dim() == 4, op, ": bad tensor ranks");
  TORCH_CHECK(key.stride(2) == 1 && key.stride(1) == key.size(2) && value.stride(2) == 1 && value.stride(1) == value.size(2), op,
              ": key/value heads must be contiguous");
  TORCH_CHECK(key_cache.stride(3) == 1 && value_cache.stride(3) == 1, op, ": cache head rows must be contiguous");
  // reference: TORCH_CHECK(key_cache.stride(0) == value_cache.stride(0)) (cache_kernels.cu:438)
  TORCH_CHECK(key_cache.stride(0) == value_cache.stride(0) && key_cache.stride(1) == value_cache.stride(1) &&
                  key_cache.stride(2) == value_cache.stride(2),
              op, ": key_cache and value_cache must share their strides");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(), op,
              ": slot_mapping must be contiguous int64");
  TORCH_CHECK(k_scale.scalar_type() == at::ScalarType::Float && v_scale.scalar_type() == at::ScalarType::Float, op,
              ": k_scale / v_scale must be float32 tensors");
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  // reference: num_tokens = slot_mapping.size(0) - with padding key.size(0) can be larger (cache_kernels.cu:423-431)
  const int st = aph_reshape_and_cache_flash(
      key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr<int64_t>(),
      slot_mapping.size(0), key.stride(0), value.stride(0), key_cache.stride(0), key_cache.stride(1), key_cache.stride(2),
      (int)key.size(1), (int)key.size(2), (int)key_cache.size(1), dtype_code(key, op), kv_code(kv_cache_dtype),
      k_scale.data_ptr<float>(), v_scale.data_ptr<float>(), stream_of());
  check_status(st, op);
}

void copy_blocks(std::vector<torch::Tensor> const& key_caches, std::vector<torch::Tensor> const& value_caches,
                 const torch::Tensor& block_mapping) {
  const char* op = "copy_blocks";
  const int64_t num_layers = (int64_t)key_caches.size();
  TORCH_CHECK(num_layers == (int64_t)value_caches.size(), op, ": key_caches and value_caches differ in length");
  if (num_layers == 0) return;
  const torch::Device dev = key_caches[0].device();
  TORCH_CHECK(dev.is_cuda(), op, ": caches must be on a CUDA device");
  TORCH_CHECK(block_mapping.is_cuda() && block_mapping.scalar_type() == at::ScalarType::Long && block_mapping.dim() == 2 &&
                  block_mapping.size(1) == 2 && block_mapping.is_contiguous(),
              op, ": block_mapping must be a contiguous int64 [num_pairs, 2] tensor on the device");
  std::vector<int64_t> ptrs(2 * num_layers);
  for (int64_t l = 0; l < num_layers; ++l) {
    ptrs[l] = reinterpret_cast<int64_t>(key_caches[l].data_ptr());
    ptrs[num_layers + l] = reinterpret_cast<int64_t>(value_caches[l].data_ptr());
  }
  // one upload for both pointer tables (the reference uploads two, synchronously: cache_kernels.cu:148-156)
  torch::Tensor ptrs_d = torch::from_blob(ptrs.data(), {2 * num_layers}, torch::kInt64).to(dev);
  const at::cuda::OptionalCUDAGuard guard(dev);
  const int64_t block_bytes = key_caches[0][0].numel() * key_caches[0].element_size();
  const int st = aph_copy_blocks(ptrs_d.data_ptr<int64_t>(), ptrs_d.data_ptr<int64_t>() + num_layers, block_mapping.data_ptr<int64_t>(),
                                 (int)num_layers, block_mapping.size(0), block_bytes, stream_of());
  check_status(st, op);
}

// kernels/cache_kernels.cu:173-208
void copy_blocks_mla(std::vector<torch::Tensor> const& kv_caches, const torch::Tensor& block_mapping) {
  const char* op = "copy_blocks_mla";
  const int64_t num_layers = (int64_t)kv_caches.size();
  if (num_layers == 0) return;
  const torch::Device dev = kv_caches[0].device();
  TORCH_CHECK(dev.is_cuda(), "kv_cache must be on CUDA");
  TORCH_CHECK(block_mapping.is_cuda() && block_mapping.scalar_type() == at::ScalarType::Long && block_mapping.dim() == 2 &&
                  block_mapping.size(1) == 2 && block_mapping.is_contiguous(),
              op, ": block_mapping must be a contiguous int64 [num_pairs, 2] tensor on the device");
  std::vector<int64_t> ptrs(num_layers);
  for (int64_t l = 0; l < num_layers; ++l) ptrs[l] = reinterpret_cast<int64_t>(kv_caches[l].data_ptr());
  torch::Tensor ptrs_d = torch::from_blob(ptrs.data(), {num_layers}, torch::kInt64).to(dev);
  const at::cuda::OptionalCUDAGuard guard(dev);
  const int64_t block_bytes = kv_caches[0].stride(0) * kv_caches[0].element_size();  // padding included, as the reference
  check_status(aph_copy_blocks_mla(ptrs_d.data_ptr<int64_t>(), block_mapping.data_ptr<int64_t>(), (int)num_layers,
                                   block_mapping.size(0), block_bytes, stream_of()), op);
}

// kernels/cache_kernels.cu:463-500
void concat_and_cache_mla(torch::Tensor& kv_c, torch::Tensor& k_pe, torch::Tensor& kv_cache, torch::Tensor& slot_mapping,
                          const std::string& kv_cache_dtype, torch::Tensor& scale) {
  const char* op = "concat_and_cache_mla";
  TORCH_CHECK(kv_c.is_cuda() && k_pe.is_cuda() && kv_cache.is_cuda() && slot_mapping.is_cuda() && scale.is_cuda(), op,
              ": tensors must be on a CUDA device");
  TORCH_CHECK(kv_c.dim() == 2 && k_pe.dim() == 2 && kv_cache.dim() == 3, op, ": kv_c, k_pe must be 2-D and kv_cache 3-D");
  const int64_t kv_lora_rank = kv_c.size(1), pe_dim = k_pe.size(1);
  TORCH_CHECK(kv_cache.size(2) == kv_lora_rank + pe_dim);
  TORCH_CHECK(kv_c.stride(1) == 1 && k_pe.stride(1) == 1 && kv_cache.stride(2) == 1, op, ": the last dimension must be contiguous");
  TORCH_CHECK(k_pe.scalar_type() == kv_c.scalar_type(), op, ": dtype mismatch");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(), op,
              ": slot_mapping must be contiguous int64");
  TORCH_CHECK(scale.scalar_type() == at::ScalarType::Float, op, ": scale must be float32");
  const int kv = kv_code(kv_cache_dtype);
  TORCH_CHECK(kv == APH_KV_AUTO ? kv_cache.scalar_type() == kv_c.scalar_type() : kv_cache.element_size() == 1, op,
              ": kv_cache dtype does not match kv_cache_dtype");
  const at::cuda::OptionalCUDAGuard guard(device_of(kv_c));
  check_status(aph_concat_and_cache_mla(kv_c.data_ptr(), k_pe.data_ptr(), kv_cache.data_ptr(), slot_mapping.data_ptr<int64_t>(),
                                        slot_mapping.size(0), (int)kv_lora_rank, (int)pe_dim, (int)kv_cache.size(1),
                                        kv_cache.stride(0), kv_cache.stride(1), kv_c.stride(0), k_pe.stride(0),
                                        dtype_code(kv_c, op), kv, scale.data_ptr<float>(), stream_of()), op);
}

// kernels/cache_kernels.cu:676-735
void gather_cache(torch::Tensor const& src_cache, torch::Tensor const& dst, torch::Tensor const& block_table,
                  torch::Tensor const& cu_seq_lens, int64_t batch_size, std::optional<torch::Tensor> seq_starts) {
  const char* op = "gather_cache";
  TORCH_CHECK(src_cache.is_cuda(), op, ": src_cache must be on a CUDA device");
  TORCH_CHECK(block_table.scalar_type() == at::ScalarType::Int, "block_table must be int32");
  TORCH_CHECK(cu_seq_lens.scalar_type() == at::ScalarType::Int, "cu_seq_lens must be int32");
  if (seq_starts.has_value()) {
    TORCH_CHECK(seq_starts->scalar_type() == at::ScalarType::Int, "seq_starts must be int32");
    TORCH_CHECK(src_cache.device() == seq_starts->device(), "src_cache and seq_starts must be on the same device");
  }
  TORCH_CHECK(src_cache.device() == dst.device(), "src_cache and dst must be on the same device");
  TORCH_CHECK(src_cache.device() == block_table.device(), "src_cache and block_table must be on the same device");
  TORCH_CHECK(src_cache.device() == cu_seq_lens.device(), "src_cache and cu_seq_lens must be on the same device");
  TORCH_CHECK(src_cache.scalar_type() == dst.scalar_type(), "src_cache and dst must have the same dtype");
  TORCH_CHECK(src_cache.dim() >= 3 && cu_seq_lens.numel() >= batch_size + 1 && cu_seq_lens.is_contiguous(), op, ": bad shapes");
  const int64_t es = src_cache.element_size();
  const int64_t entry = src_cache.flatten(2, -1).size(2);
  const at::cuda::OptionalCUDAGuard guard(device_of(src_cache));
  check_status(aph_gather_cache(src_cache.data_ptr(), dst.data_ptr(), block_table.data_ptr<int32_t>(), cu_seq_lens.data_ptr<int32_t>(),
                                seq_starts.has_value() ? seq_starts->data_ptr<int32_t>() : nullptr, (int)batch_size,
                                (int)src_cache.size(1), entry * es, block_table.stride(0), src_cache.stride(0) * es,
                                src_cache.stride(1) * es, dst.stride(0) * es, stream_of()), op);
}

void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const torch::Tensor& block_mapping) {
  const char* op = "swap_blocks";
  int kind;
  if (src.is_cuda() && dst.is_cuda()) {
    TORCH_CHECK(src.device().index() == dst.device().index(), "src and dst must be on the same GPU");
    kind = 0;
  } else if (src.is_cuda() && dst.is_cpu()) {
    kind = 1;
  } else if (src.is_cpu() && dst.is_cuda()) {
    kind = 2;
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  TORCH_CHECK(block_mapping.device().is_cpu(), "block_mapping must be on CPU");
  TORCH_CHECK(block_mapping.scalar_type() == at::ScalarType::Long && block_mapping.dim() == 2 && block_mapping.size(1) == 2, op,
              ": block_mapping must be int64 [num_pairs, 2]");
  const torch::Tensor bm = block_mapping.contiguous();
  const at::cuda::OptionalCUDAGuard guard(src.is_cuda() ? src.device() : dst.device());
  // the stride, not numel: the blocks (padding included) are contiguous (cache_kernels.cu:50-53)
  const int64_t block_bytes = src.element_size() * src.stride(0);
  const int st = aph_swap_blocks(src.data_ptr(), dst.data_ptr(), bm.data_ptr<int64_t>(), bm.size(0), block_bytes, kind, stream_of());
  check_status(st, op);
}

void check_gemm_args(const torch::Tensor& a, const torch::Tensor& b_q_weight, const torch::Tensor& b_scales,
                     const torch::Tensor& workspace, int64_t size_m, int64_t size_n, int64_t size_k) {
  const char* op = "gptq_marlin_gemm";
  TORCH_CHECK(a.is_cuda() && b_q_weight.is_cuda() && b_scales.is_cuda() && workspace.is_cuda(), op,
              ": tensors must be on a CUDA device");
  TORCH_CHECK(a.dim() == 2 && a.size(0) == size_m, "Shape mismatch: a.size(0) = ", a.size(0), ", size_m = ", size_m);
  TORCH_CHECK(a.size(1) == size_k, "Shape mismatch: a.size(1) = ", a.size(1), ", size_k = ", size_k);
  TORCH_CHECK(a.stride(1) == 1, "A must be contiguous in its last dimension");
  TORCH_CHECK(b_q_weight.is_contiguous(), "b_q_weight is not contiguous");
  TORCH_CHECK(b_scales.is_contiguous() && b_scales.dim() == 2, "b_scales must be a contiguous 2-D tensor");
  TORCH_CHECK(b_scales.scalar_type() == a.scalar_type(), "b_scales type must match a");
  TORCH_CHECK(b_scales.size(1) == size_n, "b_scales dim 1 = ", b_scales.size(1), " is not size_n = ", size_n);
  TORCH_CHECK(b_q_weight.scalar_type() == at::ScalarType::Int && workspace.scalar_type() == at::ScalarType::Int, op,
              ": b_q_weight and workspace must be int32");
}

torch::Tensor gptq_marlin_gemm(torch::Tensor& a, torch::Tensor& b_q_weight, torch::Tensor& b_scales,
                               torch::Tensor& b_zeros, torch::Tensor& g_idx, torch::Tensor& perm,
                               torch::Tensor& workspace, int64_t const& b_q_type_id, int64_t size_m, int64_t size_n,
                               int64_t size_k, bool is_k_full, bool has_zp, bool use_atomic_add, bool use_fp32_reduce,
                               bool is_zp_float) {
  const char* op = "gptq_marlin_gemm";
  check_gemm_args(a, b_q_weight, b_scales, workspace, size_m, size_n, size_k);
  const at::cuda::OptionalCUDAGuard guard(device_of(a));
  const int dt = dtype_code(a, op);
  torch::Tensor c = torch::empty({size_m, size_n}, a.options());
  if (size_m == 0) return c;
  const size_t sbytes = aph_gptq_marlin_gemm_scratch_bytes(size_m, size_n, size_k);
  torch::Tensor scratch;
  void* sptr = nullptr;
  if (sbytes > 0) {
    scratch = torch::empty({(int64_t)sbytes}, torch::TensorOptions().dtype(torch::kUInt8).device(a.device()));
    sptr = scratch.data_ptr();
  }
  const void* zp = (has_zp && b_zeros.numel() > 0) ? b_zeros.data_ptr() : nullptr;
  // Act-order (gptq_marlin.cu:2344-2400,2187-2202): g_idx and perm both of length size_k. With a full K the weight rows are
  // sorted by group, so the GEMM on the column-permuted activations is the plain grouped one.
  TORCH_CHECK((g_idx.numel() == 0 && perm.numel() == 0) || (g_idx.numel() == size_k && perm.numel() == size_k),
              "Unexpected g_idx.size(0) = ", g_idx.numel(), " and perm.size(0) = ", perm.numel(), ", where size_k = ", size_k);
  const void* a_ptr = a.data_ptr();
  int64_t lda = a.stride(0);
  torch::Tensor a_tmp;
  if (g_idx.numel() > 0) {
    TORCH_CHECK(b_scales.size(0) > 1, "For act_order, num_groups must be > 1");
    TORCH_CHECK(g_idx.is_cuda() && g_idx.scalar_type() == at::ScalarType::Int && g_idx.is_contiguous(), op,
                ": g_idx must be a contiguous int32 tensor on the device");
    TORCH_CHECK(perm.is_cuda() && perm.scalar_type() == at::ScalarType::Int && perm.is_contiguous(), op,
                ": perm must be a contiguous int32 tensor on the device");
    a_tmp = torch::empty({size_m, size_k}, a.options());
    check_status(aph_permute_cols(a_tmp.data_ptr(), a.data_ptr(), perm.data_ptr<int32_t>(), size_m, size_k, a.stride(0), stream_of()),
                 op);
    a_ptr = a_tmp.data_ptr();
    lda = size_k;
  }
  // rows sorted by group AND the whole K here (is_k_full): the plain grouped kernels serve it. A row-parallel shard
  // (is_k_full = False) keeps its per-row group index: the act-order kernel (marlin_actorder.cuh).
  const int32_t* g_ptr = (g_idx.numel() > 0 && !is_k_full) ? g_idx.data_ptr<int32_t>() : nullptr;
  const int st = aph_gptq_marlin_gemm(c.data_ptr(), a_ptr, b_q_weight.data_ptr(), b_scales.data_ptr(), zp, g_ptr,
                                      nullptr, workspace.data_ptr<int32_t>(), workspace.numel(), b_q_type_id, size_m,
                                      size_n, size_k, lda, b_scales.size(0), is_k_full, has_zp,
                                      use_atomic_add, use_fp32_reduce, is_zp_float, dt, sptr, sbytes, stream_of());
  check_status(st, op);
  return c;
}

// kernels/quantization/gptq_marlin/gptq_marlin_repack.cu:236-322 (same checks, same output shape)
torch::Tensor gptq_marlin_repack(torch::Tensor& b_q_weight, torch::Tensor& perm, int64_t size_k, int64_t size_n,
                                 int64_t num_bits) {
  const char* op = "gptq_marlin_repack";
  TORCH_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = ", num_bits);
  const int64_t pack = 32 / num_bits;
  TORCH_CHECK(size_k % 16 == 0, "size_k = ", size_k, " is not divisible by tile_k_size = ", 16);
  TORCH_CHECK(size_n % 64 == 0, "size_n = ", size_n, " is not divisible by tile_n_size = ", 64);
  TORCH_CHECK(b_q_weight.dim() == 2 && b_q_weight.size(0) == size_k / pack, "Shape mismatch: b_q_weight.size(0) = ",
              b_q_weight.size(0), ", size_k = ", size_k, ", pack_factor = ", pack);
  TORCH_CHECK(b_q_weight.size(1) == size_n, "b_q_weight.size(1) = ", b_q_weight.size(1), " is not size_n = ", size_n);
  TORCH_CHECK(b_q_weight.is_cuda(), "b_q_weight is not on GPU");
  TORCH_CHECK(b_q_weight.is_contiguous(), "b_q_weight is not contiguous");
  TORCH_CHECK(b_q_weight.scalar_type() == at::ScalarType::Int, "expected b_q_weight to be type int");
  TORCH_CHECK(perm.is_cuda(), "perm is not on GPU");
  TORCH_CHECK(perm.is_contiguous(), "perm is not contiguous");
  TORCH_CHECK(perm.scalar_type() == at::ScalarType::Int, "expected perm type int");
  TORCH_CHECK(perm.numel() == 0 || perm.numel() == size_k, "perm.size(0) = ", perm.numel(), " is not size_k = ", size_k);
  const at::cuda::OptionalCUDAGuard guard(device_of(b_q_weight));
  torch::Tensor out = torch::empty({size_k / 16, size_n * 16 / pack}, b_q_weight.options());
  const int32_t* pm = perm.numel() > 0 ? perm.data_ptr<int32_t>() : nullptr;
  check_status(aph_gptq_marlin_repack(out.data_ptr<int32_t>(), b_q_weight.data_ptr<int32_t>(), pm, size_k, size_n,
                                      (int)num_bits, stream_of()), op);
  return out;
}

// kernels/quantization/gptq_marlin/awq_marlin_repack.cu:189-268
torch::Tensor awq_marlin_repack(torch::Tensor& b_q_weight, int64_t size_k, int64_t size_n, int64_t num_bits) {
  const char* op = "awq_marlin_repack";
  TORCH_CHECK(num_bits == 4 || num_bits == 8, "num_bits must be 4 or 8. Got = ", num_bits);
  const int64_t pack = 32 / num_bits;
  TORCH_CHECK(size_k % 16 == 0, "size_k = ", size_k, " is not divisible by tile_k_size = ", 16);
  TORCH_CHECK(size_n % 64 == 0, "size_n = ", size_n, " is not divisible by tile_n_size = ", 64);
  TORCH_CHECK(b_q_weight.dim() == 2 && b_q_weight.size(0) == size_k, "b_q_weight.size(0) = ", b_q_weight.size(0),
              " is not size_k = ", size_k);
  TORCH_CHECK(b_q_weight.size(1) == size_n / pack, "Shape mismatch: b_q_weight.size(1) = ", b_q_weight.size(1),
              ", size_n = ", size_n, ", pack_factor = ", pack);
  TORCH_CHECK(b_q_weight.is_cuda(), "b_q_weight is not on GPU");
  TORCH_CHECK(b_q_weight.is_contiguous(), "b_q_weight is not contiguous");
  TORCH_CHECK(b_q_weight.scalar_type() == at::ScalarType::Int, "expected b_q_weight to be type int");
  const at::cuda::OptionalCUDAGuard guard(device_of(b_q_weight));
  torch::Tensor out = torch::empty({size_k / 16, size_n * 16 / pack}, b_q_weight.options());
  check_status(aph_awq_marlin_repack(out.data_ptr<int32_t>(), b_q_weight.data_ptr<int32_t>(), size_k, size_n,
                                     (int)num_bits, stream_of()), op);
  return out;
}

// kernels/layernorm_kernels.cu:139-154
void rms_norm(torch::Tensor& out, torch::Tensor& input, torch::Tensor& weight, double epsilon) {
  const char* op = "rms_norm";
  TORCH_CHECK(out.is_cuda() && input.is_cuda() && weight.is_cuda(), op, ": tensors must be on a CUDA device");
  TORCH_CHECK(out.is_contiguous() && input.is_contiguous() && weight.is_contiguous(), op, ": tensors must be contiguous");
  TORCH_CHECK(out.scalar_type() == input.scalar_type() && weight.scalar_type() == input.scalar_type(), op, ": dtype mismatch");
  const int64_t hidden = input.size(-1);
  TORCH_CHECK(weight.numel() == hidden && out.numel() == input.numel(), op, ": shape mismatch");
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  check_status(aph_rms_norm(out.data_ptr(), input.data_ptr(), weight.data_ptr(), (float)epsilon, input.numel() / hidden, hidden,
                            hidden, dtype_code(input, op), stream_of()), op);
}

// kernels/layernorm_kernels.cu:168-240
void fused_add_rms_norm(torch::Tensor& input, torch::Tensor& residual, torch::Tensor& weight, double epsilon) {
  const char* op = "fused_add_rms_norm";
  TORCH_CHECK(input.is_cuda() && residual.is_cuda() && weight.is_cuda(), op, ": tensors must be on a CUDA device");
  TORCH_CHECK(input.is_contiguous() && residual.is_contiguous() && weight.is_contiguous(), op, ": tensors must be contiguous");
  TORCH_CHECK(residual.scalar_type() == input.scalar_type() && weight.scalar_type() == input.scalar_type(), op, ": dtype mismatch");
  const int64_t hidden = input.size(-1);
  TORCH_CHECK(weight.numel() == hidden && residual.numel() == input.numel(), op, ": shape mismatch");
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  check_status(aph_fused_add_rms_norm(input.data_ptr(), residual.data_ptr(), weight.data_ptr(), (float)epsilon,
                                      input.numel() / hidden, hidden, dtype_code(input, op), stream_of()), op);
}

// kernels/activation_kernels.cu:64-149: out [..., d] = act(input[..., :d]) * input[..., d:]
void act_and_mul(const char* op, int act, torch::Tensor& out, torch::Tensor& input, double param) {
  TORCH_CHECK(out.is_cuda() && input.is_cuda(), op, ": tensors must be on a CUDA device");
  TORCH_CHECK(out.is_contiguous() && input.is_contiguous(), op, ": tensors must be contiguous");
  TORCH_CHECK(out.scalar_type() == input.scalar_type(), op, ": dtype mismatch");
  TORCH_CHECK(input.size(-1) % 2 == 0, op, ": the last dimension of input must be even");
  const int64_t d = input.size(-1) / 2;
  TORCH_CHECK(out.numel() * 2 == input.numel() && out.size(-1) == d, op, ": shape mismatch");
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  const int64_t tokens = d > 0 ? out.numel() / d : 0;
  // the grid's y dimension carries the tokens: long prefills go in slices of 65535 tokens
  for (int64_t t0 = 0; t0 < tokens; t0 += 65535) {
    const int64_t n = std::min<int64_t>(65535, tokens - t0);
    check_status(aph_act_and_mul(act, (char*)out.data_ptr() + t0 * d * out.element_size(),
                                 (const char*)input.data_ptr() + t0 * 2 * d * input.element_size(), n, d, (float)param,
                                 dtype_code(input, op), stream_of()), op);
  }
}
void silu_and_mul(torch::Tensor& out, torch::Tensor& input) { act_and_mul("silu_and_mul", APH_ACT_SILU_AND_MUL, out, input, 0); }
void mul_and_silu(torch::Tensor& out, torch::Tensor& input) { act_and_mul("mul_and_silu", APH_ACT_MUL_AND_SILU, out, input, 0); }
void gelu_and_mul(torch::Tensor& out, torch::Tensor& input) { act_and_mul("gelu_and_mul", APH_ACT_GELU_AND_MUL, out, input, 0); }
void gelu_tanh_and_mul(torch::Tensor& out, torch::Tensor& input) {
  act_and_mul("gelu_tanh_and_mul", APH_ACT_GELU_TANH_AND_MUL, out, input, 0);
}
void fatrelu_and_mul(torch::Tensor& out, torch::Tensor& input, double threshold) {
  act_and_mul("fatrelu_and_mul", APH_ACT_FATRELU_AND_MUL, out, input, threshold);
}

// kernels/activation_kernels.cu:151-225: element-wise
void activation(const char* op, int act, torch::Tensor& out, torch::Tensor& input) {
  TORCH_CHECK(out.is_cuda() && input.is_cuda(), op, ": tensors must be on a CUDA device");
  TORCH_CHECK(out.is_contiguous() && input.is_contiguous(), op, ": tensors must be contiguous");
  TORCH_CHECK(out.scalar_type() == input.scalar_type() && out.numel() == input.numel(), op, ": dtype or shape mismatch");
  const at::cuda::OptionalCUDAGuard guard(device_of(input));
  check_status(aph_activation(act, out.data_ptr(), input.data_ptr(), input.numel(), dtype_code(input, op), stream_of()), op);
}
void gelu_new(torch::Tensor& out, torch::Tensor& input) { activation("gelu_new", APH_ACT_GELU_NEW, out, input); }
void gelu_fast(torch::Tensor& out, torch::Tensor& input) { activation("gelu_fast", APH_ACT_GELU_FAST, out, input); }
void gelu_quick(torch::Tensor& out, torch::Tensor& input) { activation("gelu_quick", APH_ACT_GELU_QUICK, out, input); }

// kernels/pos_encoding_kernels.cu:124-196 (same shape checks)
void rotary_impl(const char* op, torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int64_t head_size,
                 torch::Tensor& cos_sin_cache, bool is_neox, int64_t rot_dim, const torch::Tensor* offsets) {
  TORCH_CHECK(positions.is_cuda() && query.is_cuda() && key.is_cuda() && cos_sin_cache.is_cuda(), op,
              ": tensors must be on a CUDA device");
  const int64_t num_tokens = offsets ? offsets->size(0) : positions.numel();
  const int64_t* off_ptr = nullptr;
  if (offsets) {
    TORCH_CHECK(positions.size(0) == num_tokens || positions.numel() == num_tokens,
                "positions must have the same num_tokens or batch_size as cos_sin_cache_offsets");
    TORCH_CHECK(offsets->is_cuda() && offsets->scalar_type() == at::ScalarType::Long && offsets->is_contiguous(), op,
                ": cos_sin_cache_offsets must be contiguous int64 on the GPU");
    TORCH_CHECK(rot_dim == cos_sin_cache.size(1), op, ": rot_dim must equal cos_sin_cache.size(1)");
    off_ptr = offsets->data_ptr<int64_t>();
  }
  const int positions_ndim = positions.dim();
  TORCH_CHECK(positions_ndim == 1 || positions_ndim == 2, "positions must have shape [num_tokens] or [batch_size, seq_len]");
  if (positions_ndim == 1) {
    TORCH_CHECK(query.size(0) == positions.size(0) && key.size(0) == positions.size(0),
                "query, key and positions must have the same number of tokens");
  } else {
    TORCH_CHECK(query.size(0) == positions.size(0) && key.size(0) == positions.size(0) && query.size(1) == positions.size(1) &&
                    key.size(1) == positions.size(1),
                "query, key and positions must have the same batch_size and seq_len");
  }
  TORCH_CHECK(positions.scalar_type() == at::ScalarType::Long && positions.is_contiguous(), op, ": positions must be contiguous int64");
  TORCH_CHECK(key.scalar_type() == query.scalar_type() && cos_sin_cache.scalar_type() == query.scalar_type(), op, ": dtype mismatch");
  TORCH_CHECK(cos_sin_cache.dim() == 2 && cos_sin_cache.is_contiguous(), op, ": cos_sin_cache must be [max_position, rot_dim]");
  if (num_tokens == 0) return;
  const int64_t query_hidden = query.numel() / num_tokens, key_hidden = key.numel() / num_tokens;
  TORCH_CHECK(query_hidden % head_size == 0);
  TORCH_CHECK(key_hidden % head_size == 0);
  const int64_t num_heads = query_hidden / head_size, num_kv_heads = key_hidden / head_size;
  TORCH_CHECK(num_heads % num_kv_heads == 0);
  const int seq_dim = positions_ndim - 1;
  // the token dimensions may be strided (q / k are slices of the qkv projection); inside a token the layout is dense
  TORCH_CHECK(query.stride(-1) == 1 && key.stride(-1) == 1, op, ": the last dimension must be contiguous");
  if (positions_ndim == 2)
    TORCH_CHECK(query.stride(0) == query.stride(1) * query.size(1) && key.stride(0) == key.stride(1) * key.size(1), op,
                ": batch and sequence dimensions must collapse into one token stride");
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  check_status(aph_rotary_embedding(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(), cos_sin_cache.data_ptr(),
                                    off_ptr, num_tokens, (int)cos_sin_cache.size(1), query.stride(seq_dim), key.stride(seq_dim),
                                    (int)num_heads, (int)num_kv_heads, (int)head_size, is_neox ? 1 : 0, dtype_code(query, op),
                                    stream_of()), op);
}
void rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int64_t head_size,
                      torch::Tensor& cos_sin_cache, bool is_neox) {
  rotary_impl("rotary_embedding", positions, query, key, head_size, cos_sin_cache, is_neox, cos_sin_cache.size(1), nullptr);
}
// kernels/pos_encoding_kernels.cu:204-290: one table holding several scalings (multi-LoRA), a row offset per token
void batched_rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int64_t head_size,
                              torch::Tensor& cos_sin_cache, bool is_neox, int64_t rot_dim, torch::Tensor& cos_sin_cache_offsets) {
  rotary_impl("batched_rotary_embedding", positions, query, key, head_size, cos_sin_cache, is_neox, rot_dim, &cos_sin_cache_offsets);
}

// kernels/prepare_inputs/advance_step.cu:68-93,190-232 (same tensor checks)
void advance_step_flashattn(int64_t num_seqs, int64_t num_queries, int64_t block_size, torch::Tensor& input_tokens,
                            torch::Tensor& sampled_token_ids, torch::Tensor& input_positions, torch::Tensor& seq_lens,
                            torch::Tensor& slot_mapping, torch::Tensor& block_tables) {
  const char* op = "advance_step_flashattn";
  auto verify = [&](const char* name, const torch::Tensor& t, int64_t s0, int64_t s1, at::ScalarType type) {
    const bool ok = t.is_cuda() && (s0 == -1 || t.size(0) == s0) && (s1 == -1 || (t.dim() > 1 && t.size(1) == s1)) &&
                    t.is_contiguous() && t.scalar_type() == type;
    TORCH_CHECK(ok, "tensor: name = ", name, ", shape = ", t.sizes(), " is_cont = ", t.is_contiguous(), ", type = ", t.dtype(),
                " is not as expected: shape = [", s0, ", ", s1, "], type = ", type);
  };
  verify("input_tokens", input_tokens, num_seqs, -1, at::kLong);
  verify("sampled_token_ids", sampled_token_ids, num_queries, 1, at::kLong);
  verify("input_positions", input_positions, num_seqs, -1, at::kLong);
  verify("seq_lens", seq_lens, num_seqs, -1, at::kInt);
  verify("slot_mapping", slot_mapping, num_seqs, -1, at::kLong);
  verify("block_tables", block_tables, num_seqs, -1, at::kInt);
  const at::cuda::OptionalCUDAGuard guard(device_of(sampled_token_ids));
  check_status(aph_advance_step_flashattn((int)num_seqs, (int)num_queries, (int)block_size, input_tokens.data_ptr<int64_t>(),
                                          sampled_token_ids.data_ptr<int64_t>(), input_positions.data_ptr<int64_t>(),
                                          seq_lens.data_ptr<int32_t>(), slot_mapping.data_ptr<int64_t>(),
                                          block_tables.data_ptr<int32_t>(), block_tables.stride(0), stream_of()), op);
}

// kernels/prepare_inputs/advance_step.cu:234-311 (same tensor checks; sampled_token_ids is read flat, as there)
void advance_step_flashinfer(int64_t num_seqs, int64_t num_queries, int64_t block_size, torch::Tensor& input_tokens,
                             torch::Tensor& sampled_token_ids, torch::Tensor& input_positions, torch::Tensor& seq_lens,
                             torch::Tensor& slot_mapping, torch::Tensor& block_tables, torch::Tensor& paged_kv_indices,
                             torch::Tensor& paged_kv_indptr, torch::Tensor& paged_kv_last_page_len,
                             torch::Tensor& block_table_bound) {
  const char* op = "advance_step_flashinfer";
  auto verify = [&](const char* name, const torch::Tensor& t, int64_t s0, at::ScalarType type) {
    const bool ok = t.is_cuda() && (s0 == -1 || (t.dim() > 0 && t.size(0) == s0)) && t.is_contiguous() && t.scalar_type() == type;
    TORCH_CHECK(ok, "tensor: name = ", name, ", shape = ", t.sizes(), " is_cont = ", t.is_contiguous(), ", type = ", t.dtype(),
                " is not as expected: shape = [", s0, ", -1], type = ", type);
  };
  verify("input_tokens", input_tokens, num_seqs, at::kLong);
  verify("sampled_token_ids", sampled_token_ids, -1, at::kLong);
  TORCH_CHECK(sampled_token_ids.numel() >= num_queries, op, ": sampled_token_ids holds fewer than num_queries = ", num_queries, " ids");
  verify("input_positions", input_positions, num_seqs, at::kLong);
  verify("seq_lens", seq_lens, num_seqs, at::kInt);
  verify("slot_mapping", slot_mapping, num_seqs, at::kLong);
  verify("block_tables", block_tables, num_seqs, at::kInt);
  verify("paged_kv_indices", paged_kv_indices, -1, at::kInt);
  verify("paged_kv_indptr", paged_kv_indptr, num_seqs + 1, at::kInt);
  verify("paged_kv_last_page_len", paged_kv_last_page_len, num_seqs, at::kInt);
  verify("block_table_bound", block_table_bound, num_seqs, at::kInt);
  const at::cuda::OptionalCUDAGuard guard(device_of(sampled_token_ids));
  check_status(aph_advance_step_flashinfer(
                   (int)num_seqs, (int)num_queries, (int)block_size, input_tokens.data_ptr<int64_t>(), sampled_token_ids.data_ptr<int64_t>(),
                   input_positions.data_ptr<int64_t>(), seq_lens.data_ptr<int32_t>(), slot_mapping.data_ptr<int64_t>(),
                   block_tables.data_ptr<int32_t>(), block_tables.dim() > 1 ? block_tables.stride(0) : 0, paged_kv_indices.data_ptr<int32_t>(),
                   paged_kv_indices.numel(), paged_kv_indptr.data_ptr<int32_t>(), paged_kv_last_page_len.data_ptr<int32_t>(),
                   block_table_bound.data_ptr<int32_t>(), stream_of()),
               op);
}

}  // namespace

#ifdef APH_DEFINE_SCHEMAS
TORCH_LIBRARY_FRAGMENT(_C, ops) {
  ops.def(
      "paged_attention_v1("
      "    Tensor! out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
      "    int tp_rank, int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.def(
      "paged_attention_v2("
      "    Tensor! out, Tensor! exp_sums, Tensor! max_logits,"
      "    Tensor! tmp_out, Tensor query, Tensor key_cache,"
      "    Tensor value_cache, int num_kv_heads, float scale,"
      "    Tensor block_tables, Tensor seq_lens, int block_size,"
      "    int max_seq_len, Tensor? alibi_slopes,"
      "    str kv_cache_dtype, Tensor k_scale, Tensor v_scale,"
      "    int tp_rank, int blocksparse_local_blocks,"
      "    int blocksparse_vert_stride, int blocksparse_block_size,"
      "    int blocksparse_head_sliding_step) -> ()");
  ops.def(
      "gptq_marlin_gemm(Tensor a, Tensor b_q_weight, Tensor b_scales, "
      "Tensor b_zeros, Tensor g_idx, Tensor perm, Tensor workspace, "
      "int b_q_type, "
      "SymInt size_m, SymInt size_n, SymInt size_k, bool is_k_full, "
      "bool has_zp, bool use_atomic_add, bool use_fp32_reduce, "
      "bool is_zp_float) -> Tensor",
      {at::Tag::needs_fixed_stride_order});
  // kernels/torch_bindings.cpp:337-347
  ops.def(
      "gptq_marlin_repack(Tensor b_q_weight, Tensor perm, "
      "SymInt size_k, SymInt size_n, int num_bits) -> Tensor");
  ops.def(
      "awq_marlin_repack(Tensor b_q_weight, SymInt size_k, "
      "SymInt size_n, int num_bits) -> Tensor");
  // kernels/torch_bindings.cpp:84,140-148,178-181
  ops.def("silu_and_mul(Tensor! result, Tensor input) -> ()");
  ops.def(
      "advance_step_flashattn(int num_seqs, int num_queries, int block_size, "
      "Tensor! input_tokens, Tensor sampled_token_ids, "
      "Tensor! input_positions, Tensor! seq_lens, Tensor! slot_mapping, "
      "Tensor block_tables) -> ()");
  ops.def(
      "advance_step_flashinfer("
      "    int num_seqs, int num_queries, int block_size,"
      "    Tensor! input_tokens, Tensor sampled_token_ids,"
      "    Tensor! input_positions, Tensor! seq_lens, Tensor! slot_mapping,"
      "    Tensor block_tables, Tensor! paged_kv_indices,"
      "    Tensor! paged_kv_indptr, Tensor! paged_kv_last_page_len,"
      "    Tensor! block_table_bounds"
      ") -> ()");
  ops.def("mul_and_silu(Tensor! out, Tensor input) -> ()");
  ops.def("gelu_and_mul(Tensor! out, Tensor input) -> ()");
  ops.def("gelu_tanh_and_mul(Tensor! out, Tensor input) -> ()");
  ops.def("fatrelu_and_mul(Tensor! out, Tensor input, float threshold) -> ()");
  ops.def("gelu_new(Tensor! out, Tensor input) -> ()");
  ops.def("gelu_fast(Tensor! out, Tensor input) -> ()");
  ops.def("gelu_quick(Tensor! out, Tensor input) -> ()");
  ops.def(
      "batched_rotary_embedding(Tensor positions, Tensor! query,"
      "                         Tensor! key, int head_size,"
      "                         Tensor cos_sin_cache, bool is_neox,"
      "                         int rot_dim,"
      "                         Tensor cos_sin_cache_offsets) -> ()");
  ops.def(
      "rms_norm(Tensor! result, Tensor input, Tensor weight, float epsilon) -> "
      "()");
  ops.def(
      "fused_add_rms_norm(Tensor! input, Tensor! residual, Tensor weight, "
      "float epsilon) -> ()");
  ops.def(
      "rotary_embedding(Tensor positions, Tensor! query,"
      "                 Tensor! key, int head_size,"
      "                 Tensor cos_sin_cache, bool is_neox) -> ()");
}
TORCH_LIBRARY_FRAGMENT(_C_cache_ops, ops) {
  ops.def(
      "reshape_and_cache(Tensor key, Tensor value,"
      "                  Tensor! key_cache, Tensor! value_cache,"
      "                  Tensor slot_mapping,"
      "                  str kv_cache_dtype,"
      "                  Tensor k_scale, Tensor v_scale) -> ()");
  ops.def(
      "convert_fp8(Tensor! dst_cache, Tensor src_cache, float scale, "
      "str kv_cache_dtype) -> ()");
  // kernels/torch_bindings.cpp:680-713
  ops.def("swap_blocks(Tensor src, Tensor! dst, Tensor block_mapping) -> ()");
  ops.def(
      "copy_blocks(Tensor(a!)[] key_caches, Tensor[](b!) value_caches, "
      "Tensor block_mapping) -> ()");
  // kernels/torch_bindings.cpp:691-693,716-722,731-734
  ops.def("copy_blocks_mla(Tensor(a!)[] kv_caches, Tensor block_mapping) -> ()");
  ops.def(
      "concat_and_cache_mla(Tensor kv_c, Tensor k_pe,"
      "                     Tensor! kv_cache,"
      "                     Tensor slot_mapping,"
      "                     str kv_cache_dtype,"
      "                     Tensor scale) -> ()");
  ops.def(
      "gather_cache(Tensor src_cache, Tensor! dst, Tensor block_table, "
      "Tensor cu_seq_lens, int batch_size, Tensor? seq_starts) -> ()");
  ops.def(
      "reshape_and_cache_flash(Tensor key, Tensor value,"
      "                        Tensor! key_cache,"
      "                        Tensor! value_cache,"
      "                        Tensor slot_mapping,"
      "                        str kv_cache_dtype,"
      "                        Tensor k_scale, Tensor v_scale) -> ()");
}
#endif

#define APH_ALL_C_OPS(X)                                                                                              \
  X(_C, paged_attention_v1) X(_C, paged_attention_v2) X(_C, gptq_marlin_gemm) X(_C, gptq_marlin_repack)                 \
  X(_C, awq_marlin_repack) X(_C, silu_and_mul) X(_C, advance_step_flashattn) X(_C, advance_step_flashinfer) X(_C, mul_and_silu) X(_C, gelu_and_mul)    \
  X(_C, gelu_tanh_and_mul) X(_C, fatrelu_and_mul) X(_C, gelu_new) X(_C, gelu_fast) X(_C, gelu_quick)                    \
  X(_C, batched_rotary_embedding) X(_C, rms_norm) X(_C, fused_add_rms_norm) X(_C, rotary_embedding)
#define APH_ALL_CACHE_OPS(X)                                                                                          \
  X(_C_cache_ops, reshape_and_cache) X(_C_cache_ops, convert_fp8) X(_C_cache_ops, swap_blocks)                        \
  X(_C_cache_ops, copy_blocks) X(_C_cache_ops, reshape_and_cache_flash) X(_C_cache_ops, copy_blocks_mla)              \
  X(_C_cache_ops, concat_and_cache_mla) X(_C_cache_ops, gather_cache)

#ifdef APH_DEFINE_SCHEMAS
// standalone: these are the CUDA kernels of the ops
#define APH_IMPL_CUDA(ns, name) ops.impl(#name, &name);
TORCH_LIBRARY_IMPL(_C, CUDA, ops) { APH_ALL_C_OPS(APH_IMPL_CUDA) }
TORCH_LIBRARY_IMPL(_C_cache_ops, CUDA, ops) { APH_ALL_CACHE_OPS(APH_IMPL_CUDA) }
#else
// Overlay onto a live aphrodite._C. The reference's CUDA kernels stay registered; ours sit on the BackendSelect key,
// which every call meets BEFORE its backend key. A call whose tensors are on a CUDA device runs here; if this library
// reports the argument combination as unsupported (nothing launched yet) - or the call is not for CUDA at all - it is
// redispatched below BackendSelect and reaches the kernel the reference registered. So a drop-in overlay never turns
// a configuration the reference serves into an error (8-bit Marlin weights, HQQ zero points, ...).
namespace {
template <const char* (*Name)(), auto Fn>
struct Overlay;
template <const char* (*Name)(), typename R, typename... A, R (*Fn)(A...)>
struct Overlay<Name, Fn> {
  static R call(c10::DispatchKeySet ks, A... args) {
    if (ks.has_backend(c10::BackendComponent::CUDABit)) {
      try {
        return Fn(args...);
      } catch (const UnsupportedHere&) {
      }
    }
    static const auto op = c10::Dispatcher::singleton().findSchemaOrThrow(Name(), "").template typed<R(A...)>();
    return op.redispatch(ks & c10::DispatchKeySet(c10::DispatchKeySet::FULL_AFTER, c10::DispatchKey::BackendSelect), args...);
  }
};
#define APH_OVERLAY_NAME(ns, name) \
  const char* aph_op_name_##name() { return #ns "::" #name; }
APH_ALL_C_OPS(APH_OVERLAY_NAME)
APH_ALL_CACHE_OPS(APH_OVERLAY_NAME)
}  // namespace
#define APH_IMPL_OVERLAY(ns, name) ops.impl(#name, &Overlay<&aph_op_name_##name, &name>::call);
TORCH_LIBRARY_IMPL(_C, BackendSelect, ops) { APH_ALL_C_OPS(APH_IMPL_OVERLAY) }
TORCH_LIBRARY_IMPL(_C_cache_ops, BackendSelect, ops) { APH_ALL_CACHE_OPS(APH_IMPL_OVERLAY) }
#endif
