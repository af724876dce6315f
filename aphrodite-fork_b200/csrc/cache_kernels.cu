// reshape_and_cache / convert_fp8 for sm_100a.
// Replaces /root/reference/kernels/cache_kernels.cu:210-262,369-395 (reshape_and_cache) and
// :499-577 (convert_fp8). Results are bit-exact with the reference (same index arithmetic,
// same __nv_cvt_float_to_fp8(x / scale, SATFINITE) codec); the design is different: the
// K side moves 128-bit input vectors into the x-chunks of the cache with one vector store each; the V
// side writes groups of 8 tokens that fill consecutive slots of a block (prefill) transposed, as one
// 8/16-byte store per cache row, and scatters lane-consecutive elements otherwise (decode); a group of
// 8 tokens is sliced over several CTAs so that decode-sized batches still spread over the SMs.
#include "common.cuh"

namespace aphb200 {

namespace {

// N elements moved as one vector. The alignment is capped at 16 bytes: that is all the callers check (aligned16), so a
// pack of 8 floats leaves as two 16-byte accesses instead of one 32-byte access that would need a 32-byte aligned cache.
template <typename T, int N>
struct alignas(sizeof(T) * N > 16 ? 16 : sizeof(T) * N) Pack {
  T v[N];
};

// Generic element-wise path (any alignment / stride): reference indexing, one element per thread.
template <typename scalar_t, typename cache_t, int KV>
__global__ void reshape_and_cache_generic_kernel(const scalar_t* __restrict__ key, const scalar_t* __restrict__ value,
                                                 cache_t* __restrict__ key_cache, cache_t* __restrict__ value_cache,
                                                 const int64_t* __restrict__ slot_mapping, int64_t key_stride,
                                                 int64_t value_stride, int num_heads, int head_size, int block_size,
                                                 int x, int64_t block_stride, const float* __restrict__ k_scale,
                                                 const float* __restrict__ v_scale) {
  pdl_trigger();
  pdl_wait();  // key / value come from the previous kernel; the cache may still be read by it
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t blk = slot / block_size, off = slot % block_size;
  const int n = num_heads * head_size;
  float ks = 1.f, vs = 1.f;
  if constexpr (KV != APH_KV_AUTO) {
    ks = *k_scale;
    vs = *v_scale;
  }
  for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < n; i += gridDim.y * blockDim.x) {
    const int h = i / head_size, e = i % head_size;
    const int64_t tk = blk * block_stride + (int64_t)h * head_size * block_size + (int64_t)(e / x) * block_size * x +
                       off * x + e % x;
    const int64_t tv = blk * block_stride + (int64_t)h * head_size * block_size + (int64_t)e * block_size + off;
    const scalar_t kv = key[token * key_stride + i];
    const scalar_t vv = value[token * value_stride + i];
    if constexpr (KV == APH_KV_AUTO) {
      key_cache[tk] = kv;
      value_cache[tv] = vv;
    } else {
      key_cache[tk] = fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(kv), ks);
      value_cache[tv] = fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(vv), vs);
    }
  }
}

// Vectorised path. One CTA serves (a slice of) a group of TG = 8 consecutive tokens; requires 16-byte aligned
// key / value rows and cache blocks.
//   K side: one work item per 128-bit input vector (LV elements); its LV converted elements are contiguous inside
//           one x-chunk of the cache ([D/x][bs][x]), so they leave as one 16 / 8 / 4-byte store.
//   V side: the layout [D][bs] gives a token ONE element in every bs-strided row, so a lone token can only scatter
//           elements. Lanes take consecutive elements (a warp store then covers one contiguous 32*bs*sizeof span,
//           the fewest cache lines this layout allows). When the TG tokens of the group occupy TG consecutive,
//           TG-aligned slots of one block (every prefill chunk does) the group is written transposed instead: one
//           thread gathers the TG tokens' values of one row and stores them as ONE 8- or 16-byte vector - 8x fewer
//           store transactions than per-element scatter.
constexpr int TG = 8;

template <typename scalar_t, typename cache_t, int KV>
__global__ void __launch_bounds__(256)
    reshape_and_cache_vec_kernel(const scalar_t* __restrict__ key, const scalar_t* __restrict__ value,
                                 cache_t* __restrict__ key_cache, cache_t* __restrict__ value_cache,
                                 const int64_t* __restrict__ slot_mapping, int64_t num_tokens, int64_t key_stride,
                                 int64_t value_stride, int num_heads, int head_size, int block_size, int64_t block_stride,
                                 const float* __restrict__ k_scale, const float* __restrict__ v_scale) {
  constexpr int X = 16 / sizeof(cache_t);        // elements per K chunk
  constexpr int LV = 16 / sizeof(scalar_t);      // elements per 128-bit input vector (LV divides X)
  using InVec = Pack<scalar_t, LV>;
  __shared__ int64_t s_slot[TG];
  __shared__ int s_run;
  pdl_trigger();
  pdl_wait();  // key / value come from the previous kernel; the cache may still be read by it
  const int64_t t0 = (int64_t)blockIdx.x * TG;
  const int nt = (int)(num_tokens - t0 < TG ? num_tokens - t0 : TG);
  if (threadIdx.x < TG) s_slot[threadIdx.x] = (int)threadIdx.x < nt ? slot_mapping[t0 + threadIdx.x] : -1;
  __syncthreads();
  if (threadIdx.x == 0) {
    bool run = nt == TG && s_slot[0] >= 0 && s_slot[0] % TG == 0 && block_size % TG == 0;
    for (int i = 1; i < TG; ++i) run = run && s_slot[i] == s_slot[0] + i;
    s_run = run;
  }
  __syncthreads();
  const int n = num_heads * head_size;
  float ks = 1.f, vs = 1.f;
  if constexpr (KV != APH_KV_AUTO) {
    ks = *k_scale;
    vs = *v_scale;
  }
  const int tid = blockIdx.y * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.y * blockDim.x;
  auto cvt = [](scalar_t v, float sc) -> cache_t {
    if constexpr (KV == APH_KV_AUTO)
      return v;
    else
      return fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(v), sc);
  };

  // ---- K: (token j, input vector ci) -> LV contiguous cache elements; two independent loads in flight
  const int vecs = n / LV;
  for (int w0 = tid; w0 < nt * vecs; w0 += 2 * nthreads) {
    InVec in[2];
    cache_t* dst[2] = {nullptr, nullptr};
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int w = w0 + u * nthreads;
      if (w >= nt * vecs) continue;
      const int j = w / vecs, ci = w - j * vecs;
      const int64_t slot = s_slot[j];
      if (slot < 0) continue;
      const int64_t blk = slot / block_size, off = slot % block_size;
      in[u] = *reinterpret_cast<const InVec*>(key + (t0 + j) * key_stride + ci * LV);
      const int i0 = ci * LV;
      const int h = i0 / head_size, e0 = i0 % head_size;
      dst[u] = key_cache + blk * block_stride + (int64_t)h * head_size * block_size + (int64_t)(e0 / X) * block_size * X +
               off * X + e0 % X;
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (dst[u] == nullptr) continue;
      Pack<cache_t, LV> o;
#pragma unroll
      for (int e = 0; e < LV; ++e) o.v[e] = cvt(in[u].v[e], ks);
      *reinterpret_cast<Pack<cache_t, LV>*>(dst[u]) = o;
    }
  }

  // ---- V
  if (s_run) {
    const int64_t blk = s_slot[0] / block_size, off = s_slot[0] % block_size;
    cache_t* vc = value_cache + blk * block_stride + off;
    const scalar_t* vrow = value + t0 * value_stride;
    for (int i = tid; i < n; i += nthreads) {
      Pack<cache_t, TG> o;
#pragma unroll
      for (int j = 0; j < TG; ++j) o.v[j] = cvt(vrow[(int64_t)j * value_stride + i], vs);
      *reinterpret_cast<Pack<cache_t, TG>*>(vc + (int64_t)i * block_size) = o;
    }
  } else {
    // four independent loads in flight per thread before the dependent stores
    const int total = nt * n;
    for (int w0 = tid; w0 < total; w0 += 4 * nthreads) {
      scalar_t v[4];
      int64_t dst[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int w = w0 + u * nthreads;
        dst[u] = -1;
        if (w < total) {
          const int j = w / n, i = w - j * n;
          const int64_t slot = s_slot[j];
          if (slot >= 0) {
            dst[u] = (slot / block_size) * block_stride + (int64_t)i * block_size + slot % block_size;
            v[u] = value[(t0 + j) * value_stride + i];
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (dst[u] >= 0) value_cache[dst[u]] = cvt(v[u], vs);
    }
  }
}

template <typename scalar_t, typename cache_t, int KV>
int launch_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                             const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                             int64_t value_stride, int num_heads, int head_size, int block_size, int x,
                             int64_t block_stride, const float* k_scale, const float* v_scale, cudaStream_t stream) {
  constexpr int X = 16 / sizeof(cache_t);
  constexpr int LV = 16 / sizeof(scalar_t);
  const int n = num_heads * head_size;
  const bool vec_ok = x == X && X % LV == 0 && head_size % X == 0 && aligned16(key) && aligned16(value) &&
                      aligned16(key_cache) && aligned16(value_cache) && key_stride % LV == 0 && value_stride % LV == 0 &&
                      (block_stride * (int64_t)sizeof(cache_t)) % 16 == 0;
  if (vec_ok) {
    const int64_t groups = (num_tokens + TG - 1) / TG;
    const int threads = 256;
    // few tokens (decode): slice each group over several CTAs until the grid covers the SMs about twice
    int ysplit = 1;
    const int sms = sm_count();
    while (groups * ysplit < 2LL * sms && (int64_t)ysplit * threads < (int64_t)TG * n) ysplit *= 2;
    // many tokens: at most 8 scattered elements per thread (memory-level parallelism comes from resident warps)
    while ((int64_t)ysplit * threads * 8 < (int64_t)TG * n && ysplit < 64) ysplit *= 2;
    dim3 grid((unsigned)groups, ysplit);
    APH_LAUNCH_PDL(reshape_and_cache_vec_kernel<scalar_t, cache_t, KV>, grid, dim3(threads), (size_t)0, stream,
                   (const scalar_t*)key, (const scalar_t*)value, (cache_t*)key_cache, (cache_t*)value_cache, slot_mapping,
                   num_tokens, key_stride, value_stride, num_heads, head_size, block_size, block_stride, k_scale, v_scale);
  } else {
    dim3 grid((unsigned)num_tokens, 1);
    const int threads = n < 512 ? ((n + 31) / 32) * 32 : 512;
    APH_LAUNCH_PDL(reshape_and_cache_generic_kernel<scalar_t, cache_t, KV>, grid, dim3(threads), (size_t)0, stream,
                   (const scalar_t*)key, (const scalar_t*)value, (cache_t*)key_cache, (cache_t*)value_cache, slot_mapping,
                   key_stride, value_stride, num_heads, head_size, block_size, x, block_stride, k_scale, v_scale);
  }
  return APH_OK;
}

template <typename scalar_t, int KV>
__global__ void convert_to_fp8_kernel(uint8_t* __restrict__ dst, const scalar_t* __restrict__ src, int64_t n,
                                      float scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(src[i]), scale);
}
template <typename scalar_t, int KV>
__global__ void convert_from_fp8_kernel(scalar_t* __restrict__ dst, const uint8_t* __restrict__ src, int64_t n,
                                        float scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = ScalarTraits<scalar_t>::from_float(fp8_decode<KV>(src[i]) * scale);
}

template <typename scalar_t>
int launch_convert(void* dst, const void* src, int64_t n, float scale, int kv, int direction, cudaStream_t stream) {
  if (n == 0) return APH_OK;
  const int threads = 256;
  const int blocks = (int)((n + threads - 1) / threads < 148 * 16 ? (n + threads - 1) / threads : 148 * 16);
  if (direction == 0) {
    if (kv == APH_KV_FP8_E5M2)
      convert_to_fp8_kernel<scalar_t, APH_KV_FP8_E5M2><<<blocks, threads, 0, stream>>>((uint8_t*)dst, (const scalar_t*)src, n, scale);
    else
      convert_to_fp8_kernel<scalar_t, APH_KV_FP8_E4M3><<<blocks, threads, 0, stream>>>((uint8_t*)dst, (const scalar_t*)src, n, scale);
  } else {
    if (kv == APH_KV_FP8_E5M2)
      convert_from_fp8_kernel<scalar_t, APH_KV_FP8_E5M2><<<blocks, threads, 0, stream>>>((scalar_t*)dst, (const uint8_t*)src, n, scale);
    else
      convert_from_fp8_kernel<scalar_t, APH_KV_FP8_E4M3><<<blocks, threads, 0, stream>>>((scalar_t*)dst, (const uint8_t*)src, n, scale);
  }
  APH_LAUNCH_CHECK();
  return APH_OK;
}

// ------------------------------------------------------------------------------------------------
// reshape_and_cache_flash (reference cache_kernels.cu:264-306,412-497): caches in the NHD layout
// [num_blocks, block_size, num_heads, head_size]. A token's head row is contiguous on both sides, so every work
// item is one 128-bit input vector of K or V -> one vector store (16 bytes, 8 for fp8 from 16-bit, 4 from f32).
// ------------------------------------------------------------------------------------------------
template <typename scalar_t, typename cache_t, int KV, bool VEC>
__global__ void __launch_bounds__(256)
    reshape_and_cache_flash_kernel(const scalar_t* __restrict__ key, const scalar_t* __restrict__ value,
                                   cache_t* __restrict__ key_cache, cache_t* __restrict__ value_cache,
                                   const int64_t* __restrict__ slot_mapping, int64_t key_stride, int64_t value_stride,
                                   int64_t block_stride, int64_t page_stride, int64_t head_stride, int num_heads,
                                   int head_size, int block_size, const float* __restrict__ k_scale,
                                   const float* __restrict__ v_scale) {
  constexpr int LV = VEC ? 16 / sizeof(scalar_t) : 1;
  pdl_trigger();
  pdl_wait();
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;
  const int64_t base = (slot / block_size) * block_stride + (slot % block_size) * page_stride;
  float ks = 1.f, vs = 1.f;
  if constexpr (KV != APH_KV_AUTO) {
    ks = *k_scale;
    vs = *v_scale;
  }
  const int vec_per_head = head_size / LV;
  const int items = num_heads * vec_per_head;
  for (int w = blockIdx.y * blockDim.x + threadIdx.x; w < 2 * items; w += gridDim.y * blockDim.x) {
    const bool is_v = w >= items;
    const int i = is_v ? w - items : w;
    const int h = i / vec_per_head, e0 = (i - h * vec_per_head) * LV;
    const scalar_t* src = (is_v ? value + token * value_stride : key + token * key_stride) + h * head_size + e0;
    cache_t* dst = (is_v ? value_cache : key_cache) + base + (int64_t)h * head_stride + e0;
    const float sc = is_v ? vs : ks;
    const Pack<scalar_t, LV> in = *reinterpret_cast<const Pack<scalar_t, LV>*>(src);
    Pack<cache_t, LV> o;
#pragma unroll
    for (int e = 0; e < LV; ++e) {
      if constexpr (KV == APH_KV_AUTO)
        o.v[e] = in.v[e];
      else
        o.v[e] = fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(in.v[e]), sc);
    }
    *reinterpret_cast<Pack<cache_t, LV>*>(dst) = o;
  }
}

template <typename scalar_t, typename cache_t, int KV>
int launch_reshape_and_cache_flash(const void* key, const void* value, void* key_cache, void* value_cache,
                                   const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                                   int64_t value_stride, int64_t block_stride, int64_t page_stride, int64_t head_stride,
                                   int num_heads, int head_size, int block_size, const float* k_scale,
                                   const float* v_scale, cudaStream_t stream) {
  constexpr int LV = 16 / sizeof(scalar_t);
  constexpr int64_t CB = sizeof(cache_t);
  const bool vec_ok = head_size % LV == 0 && aligned16(key) && aligned16(value) && aligned16(key_cache) &&
                      aligned16(value_cache) && key_stride % LV == 0 && value_stride % LV == 0 &&
                      (block_stride * CB) % 16 == 0 && (page_stride * CB) % 16 == 0 && (head_stride * CB) % 16 == 0;
  const int items = 2 * num_heads * head_size / (vec_ok ? LV : 1);
  int threads = 256, ysplit = 1;
  while (threads > 32 && threads / 2 >= items) threads /= 2;
  const int sms = sm_count();
  while (num_tokens * ysplit < 2LL * sms && ysplit * threads < items) ysplit *= 2;
  dim3 grid((unsigned)num_tokens, ysplit);
  if (vec_ok)
    APH_LAUNCH_PDL(reshape_and_cache_flash_kernel<scalar_t, cache_t, KV, true>, grid, dim3(threads), (size_t)0, stream,
                   (const scalar_t*)key, (const scalar_t*)value, (cache_t*)key_cache, (cache_t*)value_cache, slot_mapping,
                   key_stride, value_stride, block_stride, page_stride, head_stride, num_heads, head_size, block_size, k_scale,
                   v_scale);
  else
    APH_LAUNCH_PDL(reshape_and_cache_flash_kernel<scalar_t, cache_t, KV, false>, grid, dim3(threads), (size_t)0, stream,
                   (const scalar_t*)key, (const scalar_t*)value, (cache_t*)key_cache, (cache_t*)value_cache, slot_mapping,
                   key_stride, value_stride, block_stride, page_stride, head_stride, num_heads, head_size, block_size, k_scale,
                   v_scale);
  return APH_OK;
}

// ------------------------------------------------------------------------------------------------
// copy_blocks (reference cache_kernels.cu:71-170): for every layer and every (src, dst) pair copy one block of the
// key cache and of the value cache. Blocks are whole multiples of 16 bytes in every supported layout: 128-bit
// copies, the block split over gridDim.z CTAs when there are few (layer, pair) units.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    copy_blocks_kernel(const int64_t* __restrict__ key_ptrs, const int64_t* __restrict__ value_ptrs,
                       const int64_t* __restrict__ block_mapping, int64_t block_bytes, int vec16) {
  pdl_trigger();
  pdl_wait();
  const int layer = blockIdx.x, pair = blockIdx.y;
  const int64_t src_blk = block_mapping[2 * pair], dst_blk = block_mapping[2 * pair + 1];
#pragma unroll
  for (int kv = 0; kv < 2; ++kv) {
    if (kv == 1 && value_ptrs == nullptr) break;  // copy_blocks_mla: one joint cache per layer
    uint8_t* base = reinterpret_cast<uint8_t*>(kv == 0 ? key_ptrs[layer] : value_ptrs[layer]);
    const uint8_t* src = base + src_blk * block_bytes;
    uint8_t* dst = base + dst_blk * block_bytes;
    if (vec16) {
      const int64_t n = block_bytes / 16;
      for (int64_t i = blockIdx.z * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.z * blockDim.x)
        reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
    } else {
      for (int64_t i = blockIdx.z * (int64_t)blockDim.x + threadIdx.x; i < block_bytes; i += (int64_t)gridDim.z * blockDim.x)
        dst[i] = src[i];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// concat_and_cache_mla (reference cache_kernels.cu:308-348,463-500): cache row of a token = [kv_c | k_pe], optionally
// FP8-encoded with one scale. One CTA per token, VEC source elements per thread and access.
// ------------------------------------------------------------------------------------------------
template <typename scalar_t, typename cache_t, int KV, int VEC>
__global__ void __launch_bounds__(128)
    concat_and_cache_mla_kernel(const scalar_t* __restrict__ kv_c, const scalar_t* __restrict__ k_pe, cache_t* __restrict__ kv_cache,
                                const int64_t* __restrict__ slot_mapping, int64_t block_stride, int64_t entry_stride,
                                int64_t kv_c_stride, int64_t k_pe_stride, int kv_lora_rank, int pe_dim, int block_size,
                                const float* __restrict__ scale) {
  pdl_trigger();
  pdl_wait();
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // padded token
  cache_t* row = kv_cache + (slot / block_size) * block_stride + (slot % block_size) * entry_stride;
  const float sc = KV == APH_KV_AUTO ? 1.f : *scale;
  const int n = (kv_lora_rank + pe_dim) / VEC;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int e = i * VEC;
    const scalar_t* src = e < kv_lora_rank ? kv_c + token * kv_c_stride + e : k_pe + token * k_pe_stride + (e - kv_lora_rank);
    const Pack<scalar_t, VEC> x = *reinterpret_cast<const Pack<scalar_t, VEC>*>(src);
    Pack<cache_t, VEC> o;
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      if constexpr (KV == APH_KV_AUTO) o.v[j] = x.v[j];
      else o.v[j] = fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(x.v[j]), sc);
    }
    *reinterpret_cast<Pack<cache_t, VEC>*>(row + e) = o;
  }
}

template <typename scalar_t, typename cache_t, int KV>
int launch_concat_mla(const void* kv_c, const void* k_pe, void* kv_cache, const int64_t* slot_mapping, int64_t num_tokens,
                      int kv_lora_rank, int pe_dim, int block_size, int64_t block_stride, int64_t entry_stride,
                      int64_t kv_c_stride, int64_t k_pe_stride, const float* scale, cudaStream_t stream) {
  constexpr int VEC = 16 / (int)sizeof(scalar_t);
  const bool vec = kv_lora_rank % VEC == 0 && pe_dim % VEC == 0 && kv_c_stride % VEC == 0 && k_pe_stride % VEC == 0 &&
                   aligned16(kv_c) && aligned16(k_pe) && aligned16(kv_cache) && (block_stride * sizeof(cache_t)) % 16 == 0 &&
                   (entry_stride * sizeof(cache_t)) % 16 == 0;
#define APH_MLA(V)                                                                                                              \
  APH_LAUNCH_PDL((concat_and_cache_mla_kernel<scalar_t, cache_t, KV, V>), dim3((unsigned)num_tokens), dim3(128), (size_t)0, stream, \
                 (const scalar_t*)kv_c, (const scalar_t*)k_pe, (cache_t*)kv_cache, slot_mapping, block_stride, entry_stride,     \
                 kv_c_stride, k_pe_stride, kv_lora_rank, pe_dim, block_size, scale)
  if (vec) APH_MLA(VEC);
  else APH_MLA(1);
#undef APH_MLA
  return APH_OK;
}

// ------------------------------------------------------------------------------------------------
// gather_cache (reference cache_kernels.cu:582-735): dst[cu_seq_lens[b] + t] = cache[block_table[b][start_block + t / bs]][t % bs]
// for t < cu_seq_lens[b+1] - cu_seq_lens[b]: the paged rows of every sequence laid out contiguously (MLA chunked
// prefill). grid (splits, batch); a warp copies one entry at a time with 16-byte (or byte) accesses.
// ------------------------------------------------------------------------------------------------
template <typename V>
__global__ void __launch_bounds__(256)
    gather_cache_kernel(const uint8_t* __restrict__ src_cache, uint8_t* __restrict__ dst, const int32_t* __restrict__ block_table,
                        const int32_t* __restrict__ cu_seq_lens, const int32_t* __restrict__ seq_starts, int block_size,
                        int entry_bytes, int64_t block_table_stride, int64_t cache_block_stride, int64_t cache_entry_stride,
                        int64_t dst_entry_stride) {
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.y;
  const int seq_start = cu_seq_lens[b], seq_len = cu_seq_lens[b + 1] - seq_start;
  const int32_t* table = block_table + b * block_table_stride + (seq_starts != nullptr ? seq_starts[b] / block_size : 0);
  const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = entry_bytes / (int)sizeof(V);
  for (int t = blockIdx.x * warps + warp; t < seq_len; t += gridDim.x * warps) {
    const V* s = reinterpret_cast<const V*>(src_cache + table[t / block_size] * cache_block_stride + (t % block_size) * cache_entry_stride);
    V* d = reinterpret_cast<V*>(dst + (int64_t)(seq_start + t) * dst_entry_stride);
    for (int i = lane; i < n; i += 32) d[i] = s[i];
  }
}

}  // namespace
}  // namespace aphb200

using namespace aphb200;

extern "C" int aph_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                                     const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                                     int64_t value_stride, int num_kv_heads, int head_size, int block_size, int x,
                                     int64_t cache_block_stride, int dtype, int kv_dtype, const float* k_scale,
                                     const float* v_scale, void* stream) {
  APH_CHECK(num_tokens >= 0 && num_kv_heads > 0 && head_size > 0 && block_size > 0 && x > 0, APH_ERR_INVALID_ARGUMENT,
            "reshape_and_cache: bad sizes");
  APH_CHECK(head_size % x == 0, APH_ERR_INVALID_ARGUMENT, "reshape_and_cache: head_size %d not a multiple of x %d",
            head_size, x);
  APH_CHECK(kv_dtype == APH_KV_AUTO || (k_scale && v_scale), APH_ERR_INVALID_ARGUMENT,
            "reshape_and_cache: fp8 kv cache needs k_scale and v_scale");
  if (num_tokens == 0) return APH_OK;
  cudaStream_t s = (cudaStream_t)stream;
#define APH_RC(SC, CT, KVC)                                                                                         \
  return launch_reshape_and_cache<SC, CT, KVC>(key, value, key_cache, value_cache, slot_mapping, num_tokens,        \
                                               key_stride, value_stride, num_kv_heads, head_size, block_size, x,    \
                                               cache_block_stride, k_scale, v_scale, s)
#define APH_RC_DT(SC)                                                  \
  switch (kv_dtype) {                                                  \
    case APH_KV_AUTO: APH_RC(SC, SC, APH_KV_AUTO);                     \
    case APH_KV_FP8_E4M3: APH_RC(SC, uint8_t, APH_KV_FP8_E4M3);        \
    case APH_KV_FP8_E5M2: APH_RC(SC, uint8_t, APH_KV_FP8_E5M2);        \
    default: break;                                                    \
  }
  switch (dtype) {
    case APH_DT_F32: APH_RC_DT(float); break;
    case APH_DT_F16: APH_RC_DT(__half); break;
    case APH_DT_BF16: APH_RC_DT(__nv_bfloat16); break;
    default: break;
  }
#undef APH_RC_DT
#undef APH_RC
  return set_error(APH_ERR_INVALID_ARGUMENT, "reshape_and_cache: unsupported dtype %d / kv cache dtype %d", dtype,
                   kv_dtype);
}

extern "C" int aph_convert_fp8(void* dst, const void* src, int64_t n, float scale, int dtype, int kv_dtype,
                               int direction, void* stream) {
  APH_CHECK(kv_dtype == APH_KV_FP8_E4M3 || kv_dtype == APH_KV_FP8_E5M2, APH_ERR_INVALID_ARGUMENT,
            "convert_fp8: kv cache dtype must be fp8_e4m3 or fp8_e5m2");
  APH_CHECK(direction == 0 || direction == 1, APH_ERR_INVALID_ARGUMENT, "convert_fp8: bad direction");
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case APH_DT_F32: return launch_convert<float>(dst, src, n, scale, kv_dtype, direction, s);
    case APH_DT_F16: return launch_convert<__half>(dst, src, n, scale, kv_dtype, direction, s);
    case APH_DT_BF16: return launch_convert<__nv_bfloat16>(dst, src, n, scale, kv_dtype, direction, s);
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "convert_fp8: unsupported dtype %d", dtype);
  }
}

extern "C" int aph_reshape_and_cache_flash(const void* key, const void* value, void* key_cache, void* value_cache,
                                           const int64_t* slot_mapping, int64_t num_tokens, int64_t key_stride,
                                           int64_t value_stride, int64_t block_stride, int64_t page_stride,
                                           int64_t head_stride, int num_heads, int head_size, int block_size, int dtype,
                                           int kv_dtype, const float* k_scale, const float* v_scale, void* stream) {
  APH_CHECK(num_tokens >= 0 && num_heads > 0 && head_size > 0 && block_size > 0, APH_ERR_INVALID_ARGUMENT,
            "reshape_and_cache_flash: bad sizes");
  APH_CHECK(kv_dtype == APH_KV_AUTO || (k_scale && v_scale), APH_ERR_INVALID_ARGUMENT,
            "reshape_and_cache_flash: fp8 kv cache needs k_scale and v_scale");
  if (num_tokens == 0) return APH_OK;
  cudaStream_t s = (cudaStream_t)stream;
#define APH_RCF(SC, CT, KVC)                                                                                             \
  return launch_reshape_and_cache_flash<SC, CT, KVC>(key, value, key_cache, value_cache, slot_mapping, num_tokens,       \
                                                     key_stride, value_stride, block_stride, page_stride, head_stride,   \
                                                     num_heads, head_size, block_size, k_scale, v_scale, s)
#define APH_RCF_DT(SC)                                                  \
  switch (kv_dtype) {                                                   \
    case APH_KV_AUTO: APH_RCF(SC, SC, APH_KV_AUTO);                     \
    case APH_KV_FP8_E4M3: APH_RCF(SC, uint8_t, APH_KV_FP8_E4M3);        \
    case APH_KV_FP8_E5M2: APH_RCF(SC, uint8_t, APH_KV_FP8_E5M2);        \
    default: break;                                                     \
  }
  switch (dtype) {
    case APH_DT_F32: APH_RCF_DT(float); break;
    case APH_DT_F16: APH_RCF_DT(__half); break;
    case APH_DT_BF16: APH_RCF_DT(__nv_bfloat16); break;
    default: break;
  }
#undef APH_RCF_DT
#undef APH_RCF
  return set_error(APH_ERR_INVALID_ARGUMENT, "reshape_and_cache_flash: unsupported dtype %d / kv cache dtype %d", dtype,
                   kv_dtype);
}

extern "C" int aph_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs, const int64_t* block_mapping,
                               int num_layers, int64_t num_pairs, int64_t block_bytes, void* stream) {
  APH_CHECK(num_layers >= 0 && num_pairs >= 0 && block_bytes > 0, APH_ERR_INVALID_ARGUMENT, "copy_blocks: bad sizes");
  APH_CHECK(num_pairs <= 65535, APH_ERR_UNSUPPORTED, "copy_blocks: more than 65535 block pairs in one call");
  if (num_layers == 0 || num_pairs == 0) return APH_OK;
  const int64_t units = (int64_t)num_layers * num_pairs;
  int zsplit = 1;
  while (units * zsplit < 2LL * sm_count() && (int64_t)zsplit * 256 * 16 * 4 < block_bytes && zsplit < 64) zsplit *= 2;
  dim3 grid((unsigned)num_layers, (unsigned)num_pairs, (unsigned)zsplit);
  APH_LAUNCH_PDL(copy_blocks_kernel, grid, dim3(256), (size_t)0, (cudaStream_t)stream, key_cache_ptrs, value_cache_ptrs,
                 block_mapping, block_bytes, (int)(block_bytes % 16 == 0));
  return APH_OK;
}

// swap_blocks (reference cache_kernels.cu:25-67): block copies between two caches (device<->device, device<->host)
// driven by a HOST list of (src, dst) pairs. Runs of pairs that are consecutive on both sides become one copy.
extern "C" int aph_swap_blocks(const void* src, void* dst, const int64_t* block_mapping_host, int64_t num_pairs,
                               int64_t block_bytes, int kind, void* stream) {
  APH_CHECK(num_pairs >= 0 && block_bytes > 0, APH_ERR_INVALID_ARGUMENT, "swap_blocks: bad sizes");
  APH_CHECK(kind == 0 || kind == 1 || kind == 2, APH_ERR_INVALID_ARGUMENT, "swap_blocks: kind must be 0 (d2d), 1 (d2h) or 2 (h2d)");
  const cudaMemcpyKind mk = kind == 0 ? cudaMemcpyDeviceToDevice : kind == 1 ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
  for (int64_t i = 0; i < num_pairs;) {
    const int64_t s0 = block_mapping_host[2 * i], d0 = block_mapping_host[2 * i + 1];
    int64_t run = 1;
    // inside ONE tensor (src == dst) a merged run could overlap itself (blocks 0,1 -> 1,2), where the reference's
    // block-by-block copies are well defined: merge only runs whose source and destination ranges are disjoint
    const bool same = kind == 0 && src == dst;
    while (i + run < num_pairs && block_mapping_host[2 * (i + run)] == s0 + run && block_mapping_host[2 * (i + run) + 1] == d0 + run &&
           (!same || (s0 + run < d0 || d0 + run < s0)))
      ++run;
    const cudaError_t e = cudaMemcpyAsync(static_cast<char*>(dst) + d0 * block_bytes, static_cast<const char*>(src) + s0 * block_bytes,
                                          (size_t)(run * block_bytes), mk, (cudaStream_t)stream);
    APH_CHECK(e == cudaSuccess, APH_ERR_CUDA, "swap_blocks: cudaMemcpyAsync failed: %s", cudaGetErrorString(e));
    i += run;
  }
  return APH_OK;
}

// copy_blocks_mla (reference cache_kernels.cu:100-117,173-208): copy_blocks over one joint cache per layer.
extern "C" int aph_copy_blocks_mla(const int64_t* kv_cache_ptrs, const int64_t* block_mapping, int num_layers, int64_t num_pairs,
                                   int64_t block_bytes, void* stream) {
  return aph_copy_blocks(kv_cache_ptrs, nullptr, block_mapping, num_layers, num_pairs, block_bytes, stream);
}

extern "C" int aph_concat_and_cache_mla(const void* kv_c, const void* k_pe, void* kv_cache, const int64_t* slot_mapping,
                                        int64_t num_tokens, int kv_lora_rank, int pe_dim, int block_size, int64_t block_stride,
                                        int64_t entry_stride, int64_t kv_c_stride, int64_t k_pe_stride, int dtype, int kv_dtype,
                                        const float* scale, void* stream) {
  const char* op = "concat_and_cache_mla";
  APH_CHECK(num_tokens >= 0 && kv_lora_rank > 0 && pe_dim >= 0 && block_size > 0, APH_ERR_INVALID_ARGUMENT, "%s: bad sizes", op);
  APH_CHECK(kv_dtype == APH_KV_AUTO || scale != nullptr, APH_ERR_INVALID_ARGUMENT, "%s: fp8 kv cache needs a scale", op);
  if (num_tokens == 0) return APH_OK;
  APH_CHECK(kv_c && k_pe && kv_cache && slot_mapping, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  cudaStream_t s = (cudaStream_t)stream;
#define APH_MLA_KV(SC, CT, KVC)                                                                                            \
  return launch_concat_mla<SC, CT, KVC>(kv_c, k_pe, kv_cache, slot_mapping, num_tokens, kv_lora_rank, pe_dim, block_size, \
                                        block_stride, entry_stride, kv_c_stride, k_pe_stride, scale, s)
#define APH_MLA_DT(SC)                                                     \
  switch (kv_dtype) {                                                      \
    case APH_KV_AUTO: APH_MLA_KV(SC, SC, APH_KV_AUTO);                     \
    case APH_KV_FP8_E4M3: APH_MLA_KV(SC, uint8_t, APH_KV_FP8_E4M3);        \
    case APH_KV_FP8_E5M2: APH_MLA_KV(SC, uint8_t, APH_KV_FP8_E5M2);        \
    default: break;                                                        \
  }
  switch (dtype) {
    case APH_DT_F32: APH_MLA_DT(float); break;
    case APH_DT_F16: APH_MLA_DT(__half); break;
    case APH_DT_BF16: APH_MLA_DT(__nv_bfloat16); break;
    default: break;
  }
#undef APH_MLA_DT
#undef APH_MLA_KV
  return set_error(APH_ERR_INVALID_ARGUMENT, "%s: unsupported dtype %d / kv cache dtype %d", op, dtype, kv_dtype);
}

extern "C" int aph_gather_cache(const void* src_cache, void* dst, const int32_t* block_table, const int32_t* cu_seq_lens,
                                const int32_t* seq_starts, int batch_size, int block_size, int64_t entry_bytes,
                                int64_t block_table_stride, int64_t cache_block_stride_bytes, int64_t cache_entry_stride_bytes,
                                int64_t dst_entry_stride_bytes, void* stream) {
  const char* op = "gather_cache";
  APH_CHECK(batch_size >= 0 && batch_size <= 65535 && block_size > 0 && entry_bytes > 0 && entry_bytes < (1LL << 31),
            APH_ERR_INVALID_ARGUMENT, "%s: bad sizes", op);
  if (batch_size == 0) return APH_OK;
  APH_CHECK(src_cache && dst && block_table && cu_seq_lens, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  // enough CTAs for two waves whatever the batch size; a CTA's 8 warps take tokens round-robin
  const int splits = std::max(1, std::min(1024, (4 * sm_count() + batch_size - 1) / batch_size));
  const dim3 grid((unsigned)splits, (unsigned)batch_size);
  const bool vec = entry_bytes % 16 == 0 && cache_block_stride_bytes % 16 == 0 && cache_entry_stride_bytes % 16 == 0 &&
                   dst_entry_stride_bytes % 16 == 0 && aligned16(src_cache) && aligned16(dst);
  if (vec)
    APH_LAUNCH_PDL(gather_cache_kernel<uint4>, grid, dim3(256), (size_t)0, (cudaStream_t)stream, (const uint8_t*)src_cache, (uint8_t*)dst,
                   block_table, cu_seq_lens, seq_starts, block_size, (int)entry_bytes, block_table_stride, cache_block_stride_bytes,
                   cache_entry_stride_bytes, dst_entry_stride_bytes);
  else
    APH_LAUNCH_PDL(gather_cache_kernel<uint8_t>, grid, dim3(256), (size_t)0, (cudaStream_t)stream, (const uint8_t*)src_cache,
                   (uint8_t*)dst, block_table, cu_seq_lens, seq_starts, block_size, (int)entry_bytes, block_table_stride,
                   cache_block_stride_bytes, cache_entry_stride_bytes, dst_entry_stride_bytes);
  return APH_OK;
}
