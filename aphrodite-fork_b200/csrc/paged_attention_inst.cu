// One translation unit per head size (compiled with -DAPH_PA_HEAD_SIZE=<D>) so that the
// template instantiations build in parallel. Defines launch_pa_d<D>().
#include <algorithm>

#include "paged_attention_kernel.cuh"

#ifndef APH_PA_HEAD_SIZE
#error "compile with -DAPH_PA_HEAD_SIZE=<head size>"
#endif

namespace aphb200 {
namespace pa {

#define APH_CAT2(a, b) a##b
#define APH_CAT(a, b) APH_CAT2(a, b)

template <typename scalar_t, int KV, int BS, bool DENSE>
static int launch_main(const Params& p, cudaStream_t stream) {
  constexpr int D = APH_PA_HEAD_SIZE;
  using C = Cfg<scalar_t, KV, D, BS, DENSE>;
  auto kern = paged_attention_kernel<scalar_t, KV, D, BS, DENSE>;
  static bool attr_set[64] = {false};  // per device; idempotent, so a benign race at worst
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES) != cudaSuccess)
      return set_error(APH_ERR_CUDA, "paged_attention: cannot reserve %d bytes of shared memory", C::SMEM_BYTES);
    attr_set[dev] = true;
  }
  const int G = p.num_heads / p.num_kv_heads;
  const int tiles = (G + HEADS_PER_CTA - 1) / HEADS_PER_CTA;
  using std::min;
  dim3 grid(p.num_kv_heads * tiles, p.num_seqs, p.num_partitions);
  if (p.cluster_parts > 1 && p.cluster_parts * ((min(G, HEADS_PER_CTA) + p.cluster_parts - 1) / p.cluster_parts) > C::RECV_SLOTS)
    return set_error(APH_ERR_INVALID_ARGUMENT, "paged_attention: cluster of %d partitions does not fit the merge buffer", p.cluster_parts);
  const cudaError_t e = launch_pdl_cluster(kern, grid, dim3(C::NTHREADS), (size_t)C::SMEM_BYTES, stream, p.cluster_parts, p);
  if (e != cudaSuccess) return set_error(APH_ERR_CUDA, "CUDA launch failed: %s", cudaGetErrorString(e));
  return APH_OK;
}

template <typename scalar_t, int KV, int BS>
static int launch_one(const Params& p, cudaStream_t stream) {
  constexpr int D = APH_PA_HEAD_SIZE;
  // the dense build exists where it buys a fourth CTA per SM (dense_variant_exists() in paged_attention.cu says the same)
  constexpr bool HAS_DENSE = KV != APH_KV_AUTO && (D == 64 || D == 128) &&
                             Cfg<scalar_t, KV, D, BS, true>::CTAS_PER_SM > Cfg<scalar_t, KV, D, BS, false>::CTAS_PER_SM;
  int st;
  if constexpr (HAS_DENSE) {
    st = p.dense ? launch_main<scalar_t, KV, BS, true>(p, stream) : launch_main<scalar_t, KV, BS, false>(p, stream);
  } else {
    st = launch_main<scalar_t, KV, BS, false>(p, stream);
  }
  if (st != APH_OK) return st;
  const int n_out = p.num_partitions / p.cluster_parts;  // partials per (sequence, head) left in the workspace
  if (n_out > 1) {
    // second level: the reduce grid is set up while the main kernel drains (it waits before reading the partials)
    Params pr = p;
    pr.num_partitions = n_out;
    pr.partition_tokens = p.partition_tokens * p.cluster_parts;
    APH_LAUNCH_PDL(paged_attention_reduce_kernel<scalar_t, D>, dim3(p.num_heads, p.num_seqs), dim3(D >= 128 ? 128 : 64), (size_t)0,
                   stream, pr, KV != APH_KV_AUTO);
  }
  return APH_OK;
}

template <typename scalar_t, int KV>
static int launch_bs(const Params& p, int block_size, cudaStream_t stream) {
  switch (block_size) {
    case 16: return launch_one<scalar_t, KV, 16>(p, stream);
#ifndef APH_FAST_BUILD
    case 8: return launch_one<scalar_t, KV, 8>(p, stream);
    case 32: return launch_one<scalar_t, KV, 32>(p, stream);
#endif
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "Unsupported block size: %d", block_size);
  }
}

int APH_CAT(launch_pa_d, APH_PA_HEAD_SIZE)(const Params& p, int dtype, int kv_dtype, int block_size,
                                          cudaStream_t stream) {
  if (kv_dtype == APH_KV_AUTO) {
    if (dtype == APH_DT_BF16) return launch_bs<__nv_bfloat16, APH_KV_AUTO>(p, block_size, stream);
    if (dtype == APH_DT_F16) return launch_bs<__half, APH_KV_AUTO>(p, block_size, stream);
    return launch_bs<float, APH_KV_AUTO>(p, block_size, stream);  // float32 cache (x = 4)
  }
  if constexpr (APH_PA_HEAD_SIZE % 16 == 0) {  // an FP8 cache stores 16 elements per chunk
    if (kv_dtype == APH_KV_FP8_E4M3) {
      if (dtype == APH_DT_BF16) return launch_bs<__nv_bfloat16, APH_KV_FP8_E4M3>(p, block_size, stream);
      if (dtype == APH_DT_F16) return launch_bs<__half, APH_KV_FP8_E4M3>(p, block_size, stream);
      return launch_bs<float, APH_KV_FP8_E4M3>(p, block_size, stream);
    }
    if (kv_dtype == APH_KV_FP8_E5M2) {
      if (dtype == APH_DT_BF16) return launch_bs<__nv_bfloat16, APH_KV_FP8_E5M2>(p, block_size, stream);
      if (dtype == APH_DT_F16) return launch_bs<__half, APH_KV_FP8_E5M2>(p, block_size, stream);
      return launch_bs<float, APH_KV_FP8_E5M2>(p, block_size, stream);
    }
  }
  return set_error(APH_ERR_INVALID_ARGUMENT, "paged_attention: unsupported kv cache dtype %d", kv_dtype);
}

}  // namespace pa
}  // namespace aphb200
