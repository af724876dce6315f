// reshape_and_cache / convert_fp8 for sm_100a.
// Replaces /root/reference/kernels/cache_kernels.cu:210-262,369-395 (reshape_and_cache) and
// :499-577 (convert_fp8). Results are bit-exact with the reference (same index arithmetic,
// same __nv_cvt_float_to_fp8(x / scale, SATFINITE) codec); the design is different: the
// K side moves whole x-element chunks (one 16-byte store per chunk, 128-bit loads), the
// V side reads 128-bit vectors and scatters, and one CTA handles a slice of a token so that
// decode-sized batches still spread over the SMs.
#include "common.cuh"

namespace aphb200 {

namespace {

template <typename T, int N>
struct alignas(sizeof(T) * N) Pack {
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

// Vectorised path. X = elements per K chunk (16 / sizeof(cache_t)); requires 16-byte aligned
// key/value rows. Work item = one K chunk (X elements -> one 16-byte store) and, for V, one
// 16-byte input vector scattered element-wise (the layout puts one element per token in every
// block_size-strided row, so the V store cannot be widened).
template <typename scalar_t, typename cache_t, int KV>
__global__ void __launch_bounds__(256)
    reshape_and_cache_vec_kernel(const scalar_t* __restrict__ key, const scalar_t* __restrict__ value,
                                 cache_t* __restrict__ key_cache, cache_t* __restrict__ value_cache,
                                 const int64_t* __restrict__ slot_mapping, int64_t key_stride, int64_t value_stride,
                                 int num_heads, int head_size, int block_size, int64_t block_stride,
                                 const float* __restrict__ k_scale, const float* __restrict__ v_scale) {
  constexpr int X = 16 / sizeof(cache_t);        // elements per K chunk
  constexpr int LV = 16 / sizeof(scalar_t);      // elements per 128-bit input vector
  using InVec = Pack<scalar_t, LV>;
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
  const scalar_t* krow = key + token * key_stride;
  const scalar_t* vrow = value + token * value_stride;
  cache_t* kc = key_cache + blk * block_stride;
  cache_t* vc = value_cache + blk * block_stride;
  const int tid = blockIdx.y * blockDim.x + threadIdx.x;
  const int nthreads = gridDim.y * blockDim.x;

  // K: chunk ci covers elements [ci*X, ci*X+X) of the token row (head_size % X == 0).
  const int chunks_per_head = head_size / X;
  for (int ci = tid; ci < n / X; ci += nthreads) {
    const int h = ci / chunks_per_head, c = ci % chunks_per_head;
    Pack<cache_t, X> o;
#pragma unroll
    for (int j = 0; j < X / LV; ++j) {
      const InVec in = *reinterpret_cast<const InVec*>(krow + ci * X + j * LV);
#pragma unroll
      for (int e = 0; e < LV; ++e) {
        if constexpr (KV == APH_KV_AUTO)
          o.v[j * LV + e] = in.v[e];
        else
          o.v[j * LV + e] = fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(in.v[e]), ks);
      }
    }
    cache_t* dst = kc + (int64_t)h * head_size * block_size + (int64_t)c * block_size * X + off * X;
    *reinterpret_cast<Pack<cache_t, X>*>(dst) = o;
  }
  // V: 128-bit loads, element scatter.
  for (int vi = tid; vi < n / LV; vi += nthreads) {
    const InVec in = *reinterpret_cast<const InVec*>(vrow + vi * LV);
    const int i0 = vi * LV;
    const int h = i0 / head_size, e0 = i0 % head_size;  // LV divides head_size
    cache_t* dst = vc + (int64_t)h * head_size * block_size + (int64_t)e0 * block_size + off;
#pragma unroll
    for (int e = 0; e < LV; ++e) {
      if constexpr (KV == APH_KV_AUTO)
        dst[(int64_t)e * block_size] = in.v[e];
      else
        dst[(int64_t)e * block_size] = fp8_encode<KV>(ScalarTraits<scalar_t>::to_float(in.v[e]), vs);
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
  const bool vec_ok = x == X && head_size % X == 0 && head_size % LV == 0 && aligned16(key) && aligned16(value) &&
                      aligned16(key_cache) && key_stride % LV == 0 && value_stride % LV == 0 &&
                      (block_stride * (int64_t)sizeof(cache_t)) % 16 == 0;
  if (vec_ok) {
    const int items = n / LV;  // V vectors dominate the item count
    int threads = 256;
    // spread one token over several CTAs when the batch is small (decode): aim for >= 2 waves
    int ysplit = 1;
    const int sms = sm_count();
    while (ysplit * 2 * threads <= items && num_tokens * ysplit < 2LL * sms) ysplit *= 2;
    while (threads > 32 && threads / 2 >= items) threads /= 2;
    dim3 grid((unsigned)num_tokens, ysplit);
    reshape_and_cache_vec_kernel<scalar_t, cache_t, KV><<<grid, threads, 0, stream>>>(
        (const scalar_t*)key, (const scalar_t*)value, (cache_t*)key_cache, (cache_t*)value_cache, slot_mapping,
        key_stride, value_stride, num_heads, head_size, block_size, block_stride, k_scale, v_scale);
  } else {
    dim3 grid((unsigned)num_tokens, 1);
    const int threads = n < 512 ? ((n + 31) / 32) * 32 : 512;
    reshape_and_cache_generic_kernel<scalar_t, cache_t, KV><<<grid, threads, 0, stream>>>(
        (const scalar_t*)key, (const scalar_t*)value, (cache_t*)key_cache, (cache_t*)value_cache, slot_mapping,
        key_stride, value_stride, num_heads, head_size, block_size, x, block_stride, k_scale, v_scale);
  }
  APH_LAUNCH_CHECK();
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
