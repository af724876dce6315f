// The glue of a decode layer for sm_100a: rms_norm, fused_add_rms_norm, silu_and_mul,
// rotary_embedding. Replaces /root/reference/kernels/layernorm_kernels.cu:15-240, activation_kernels.cu:12-91 and
// pos_encoding_kernels.cu:10-196. All four are a few hundred KiB at decode batch sizes, i.e. launch- and
// latency-bound: every thread moves 16-byte vectors, a row is read from HBM exactly once (the normalisation keeps it in
// registers between the sum of squares and the scaling; the reference reads it twice), and each kernel takes part in
// the programmatic-dependent-launch chain of the step.
//
// Arithmetic follows the reference operation by operation, including where it rounds to the storage type:
//   rms_norm            out = T(float(x) * rsqrtf(mean(x^2) + eps)) * w           (T multiply)
//   fused_add_rms_norm  z = T(x + residual) (T add); residual = z; x = T(float(z) * rs) * w
//   silu_and_mul        out = T(float(x) / (1 + expf(-float(x)))) * y              (T multiply)
//   rotary_embedding    x' = x*cos - y*sin, y' = y*cos + x*sin, every product and sum rounded to T
// so silu_and_mul and rotary_embedding are bit-identical with the reference kernels and the two norms differ only
// through the summation order of the fp32 sum of squares.
#include "common.cuh"

namespace aphb200 {
namespace {

template <typename T, int N>
struct alignas(sizeof(T) * N) Vec {
  T v[N];
};

template <typename T>
__device__ __forceinline__ T mul_t(T a, T b) {  // the reference's `scalar_t * scalar_t`: exact product, one rounding
  return ScalarTraits<T>::from_float(ScalarTraits<T>::to_float(a) * ScalarTraits<T>::to_float(b));
}
template <typename T>
__device__ __forceinline__ T add_t(T a, T b) {
  return ScalarTraits<T>::from_float(ScalarTraits<T>::to_float(a) + ScalarTraits<T>::to_float(b));
}
template <typename T>
__device__ __forceinline__ T sub_t(T a, T b) {
  return ScalarTraits<T>::from_float(ScalarTraits<T>::to_float(a) - ScalarTraits<T>::to_float(b));
}

__device__ __forceinline__ float block_sum(float v, float* s_part) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = (blockDim.x + 31) >> 5;
  if (lane == 0) s_part[warp] = v;
  __syncthreads();
  v = lane < nwarp ? s_part[lane] : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One CTA per token. VEC elements per 16-byte access (1 = unaligned / odd hidden size fallback). Up to ITER vectors
// per thread stay in registers between the two phases; longer rows re-read what phase one wrote (L2 hits).
template <typename T, int VEC, int ITER, bool FUSED_ADD, int MAXT>
__global__ void __launch_bounds__(MAXT, MAXT == 1024 ? 1 : 5)
    rms_norm_kernel(T* __restrict__ out, T* __restrict__ input, T* __restrict__ residual, const T* __restrict__ weight,
                    float epsilon, int hidden, int64_t in_stride) {
  using V = Vec<T, VEC>;
  __shared__ float s_part[32];
  const int nvec = hidden / VEC;
  const int64_t row = blockIdx.x;
  V* in_v = reinterpret_cast<V*>(input + row * in_stride);
  V* res_v = FUSED_ADD ? reinterpret_cast<V*>(residual + row * (int64_t)hidden) : nullptr;
  V* out_v = FUSED_ADD ? in_v : reinterpret_cast<V*>(out + row * (int64_t)hidden);
  const V* w_v = reinterpret_cast<const V*>(weight);
  pdl_trigger();
  // few tokens (latency-bound): the weights are immutable, fetch them before waiting on the previous kernel. Many tokens
  // (bandwidth-bound): read them when needed (8 KiB, L1-resident) and keep the registers for occupancy.
  constexpr bool PRELOAD_W = MAXT == 1024;
  V wreg[PRELOAD_W ? ITER : 1];
  if constexpr (PRELOAD_W) {
#pragma unroll
    for (int i = 0; i < ITER; ++i) {
      const int idx = threadIdx.x + i * blockDim.x;
      if (idx < nvec) wreg[i] = w_v[idx];
    }
  }
  pdl_wait();
  V keep[ITER];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < ITER; ++i) {
    const int idx = threadIdx.x + i * blockDim.x;
    if (idx < nvec) {
      V x = in_v[idx];
      if constexpr (FUSED_ADD) {
        const V r = res_v[idx];
#pragma unroll
        for (int e = 0; e < VEC; ++e) x.v[e] = add_t(x.v[e], r.v[e]);
        res_v[idx] = x;
      }
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        const float f = ScalarTraits<T>::to_float(x.v[e]);
        ss += f * f;
      }
      keep[i] = x;
    }
  }
  for (int idx = threadIdx.x + ITER * blockDim.x; idx < nvec; idx += blockDim.x) {
    V x = in_v[idx];
    if constexpr (FUSED_ADD) {
      const V r = res_v[idx];
#pragma unroll
      for (int e = 0; e < VEC; ++e) x.v[e] = add_t(x.v[e], r.v[e]);
      res_v[idx] = x;
    }
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      const float f = ScalarTraits<T>::to_float(x.v[e]);
      ss += f * f;
    }
  }
  ss = block_sum(ss, s_part);
  const float rs = rsqrtf(ss / (float)hidden + epsilon);
#pragma unroll
  for (int i = 0; i < ITER; ++i) {
    const int idx = threadIdx.x + i * blockDim.x;
    if (idx < nvec) {
      V x = keep[i];
      const V w = PRELOAD_W ? wreg[PRELOAD_W ? i : 0] : w_v[idx];
#pragma unroll
      for (int e = 0; e < VEC; ++e) x.v[e] = mul_t(ScalarTraits<T>::from_float(ScalarTraits<T>::to_float(x.v[e]) * rs), w.v[e]);
      out_v[idx] = x;
    }
  }
  for (int idx = threadIdx.x + ITER * blockDim.x; idx < nvec; idx += blockDim.x) {
    V x = FUSED_ADD ? res_v[idx] : in_v[idx];
    const V w = w_v[idx];
#pragma unroll
    for (int e = 0; e < VEC; ++e) x.v[e] = mul_t(ScalarTraits<T>::from_float(ScalarTraits<T>::to_float(x.v[e]) * rs), w.v[e]);
    out_v[idx] = x;
  }
}

// The activation functions of activation_kernels.cu:35-62,111-115,183-205, rounding to T where the reference does.
enum : int { ACT_SILU = 0, ACT_GELU = 1, ACT_GELU_TANH = 2, ACT_FATRELU = 3, ACT_GELU_NEW = 4, ACT_GELU_FAST = 5, ACT_GELU_QUICK = 6 };

template <typename T, int ACT>
__device__ __forceinline__ T act_fn(T x, float param) {
  using S = ScalarTraits<T>;
  const float f = S::to_float(x);
  if constexpr (ACT == ACT_SILU) {
    return S::from_float(__fdiv_rn(f, 1.0f + expf(-f)));
  } else if constexpr (ACT == ACT_GELU) {
    constexpr float ALPHA = M_SQRT1_2;
    return S::from_float(f * 0.5f * (1.0f + ::erf(f * ALPHA)));
  } else if constexpr (ACT == ACT_GELU_TANH) {
    constexpr float BETA = M_SQRT2 * M_2_SQRTPI * 0.5f;
    constexpr float KAPPA = 0.044715;
    float x_cube = f * f * f;
    float inner = BETA * (f + KAPPA * x_cube);
    return S::from_float(0.5f * f * (1.0f + ::tanhf(inner)));
  } else if constexpr (ACT == ACT_FATRELU) {
    return S::from_float(f > param ? f : 0.0f);
  } else if constexpr (ACT == ACT_GELU_NEW) {
    const float x3 = S::to_float(mul_t(mul_t(x, x), x));
    const T t = S::from_float(tanhf(S::to_float(S::from_float(0.79788456f * S::to_float(add_t(x, S::from_float(0.044715f * x3)))))));
    return mul_t(mul_t(S::from_float(0.5f), x), add_t(S::from_float(1.0f), t));
  } else if constexpr (ACT == ACT_GELU_FAST) {
    const T arg = mul_t(S::from_float(f * 0.79788456f), add_t(S::from_float(1.0f), mul_t(S::from_float(0.044715f * f), x)));
    const T t = S::from_float(tanhf(S::to_float(arg)));
    return mul_t(mul_t(S::from_float(0.5f), x), add_t(S::from_float(1.0f), t));
  } else {
    return S::from_float(__fdiv_rn(f, 1.0f + expf(-1.702f * f)));
  }
}

// out[t][i] = act(in[t][i]) * in[t][d + i]  (ACT_FIRST) or in[t][i] * act(in[t][d + i]).  grid (ceil(d / (VEC * 256)), tokens)
template <typename T, int VEC, int ACT, bool ACT_FIRST>
__global__ void __launch_bounds__(256)
    act_and_mul_kernel(T* __restrict__ out, const T* __restrict__ input, int d, float param) {
  using V = Vec<T, VEC>;
  pdl_trigger();
  pdl_wait();
  const int64_t token = blockIdx.y;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= d / VEC) return;
  const V x = reinterpret_cast<const V*>(input + token * 2 * d)[idx];
  const V y = reinterpret_cast<const V*>(input + token * 2 * d + d)[idx];
  V o;
#pragma unroll
  for (int e = 0; e < VEC; ++e)
    o.v[e] = ACT_FIRST ? mul_t(act_fn<T, ACT>(x.v[e], param), y.v[e]) : mul_t(x.v[e], act_fn<T, ACT>(y.v[e], param));
  reinterpret_cast<V*>(out + token * d)[idx] = o;
}

// out[i] = act(in[i]) over a flat array
template <typename T, int VEC, int ACT>
__global__ void __launch_bounds__(256) activation_kernel(T* __restrict__ out, const T* __restrict__ input, int64_t n) {
  using V = Vec<T, VEC>;
  pdl_trigger();
  pdl_wait();
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n / VEC) return;
  const V x = reinterpret_cast<const V*>(input)[idx];
  V o;
#pragma unroll
  for (int e = 0; e < VEC; ++e) o.v[e] = act_fn<T, ACT>(x.v[e], 0.f);
  reinterpret_cast<V*>(out)[idx] = o;
}

// One CTA per token; a thread owns one (head, rotation pair) at a time, query heads first, then key heads.
template <typename T, bool NEOX>
__global__ void __launch_bounds__(512)
    rotary_embedding_kernel(const int64_t* __restrict__ positions, T* __restrict__ query, T* __restrict__ key,
                            const T* __restrict__ cos_sin_cache, const int64_t* __restrict__ cache_offsets, int rot_dim,
                            int64_t query_stride, int64_t key_stride, int num_heads, int num_kv_heads, int head_size) {
  pdl_trigger();
  const int token = blockIdx.x;
  const int embed = rot_dim / 2;
  pdl_wait();  // positions may come from advance_step_flashattn earlier in the same chain: read them after the wait
  const T* cache = cos_sin_cache + (positions[token] + (cache_offsets ? cache_offsets[token] : 0)) * rot_dim;
  const int nq = num_heads * embed, total = nq + (key != nullptr ? num_kv_heads * embed : 0);
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const bool is_q = i < nq;
    const int j = is_q ? i : i - nq;
    const int head = j / embed, off = j % embed;
    T* arr = (is_q ? query + token * query_stride : key + token * key_stride) + (int64_t)head * head_size;
    const int xi = NEOX ? off : 2 * off, yi = NEOX ? embed + off : 2 * off + 1;
    const T c = cache[off], s = cache[embed + off];
    const T x = arr[xi], y = arr[yi];
    arr[xi] = sub_t(mul_t(x, c), mul_t(y, s));
    arr[yi] = add_t(mul_t(y, c), mul_t(x, s));
  }
}

// 16-byte version: a thread owns PV consecutive rotation pairs of one head. NeoX: PV = 8 (2-byte types), the x and y
// halves are two vectors; GPT-J: PV = 4, one vector holds the four interleaved (x, y) pairs.
template <typename T, bool NEOX>
__global__ void __launch_bounds__(512)
    rotary_embedding_vec_kernel(const int64_t* __restrict__ positions, T* __restrict__ query, T* __restrict__ key,
                                const T* __restrict__ cos_sin_cache, const int64_t* __restrict__ cache_offsets, int rot_dim,
                                int64_t query_stride, int64_t key_stride, int num_heads, int num_kv_heads, int head_size) {
  constexpr int VEC = 16 / (int)sizeof(T);
  constexpr int PV = NEOX ? VEC : VEC / 2;
  using V = Vec<T, VEC>;
  using H = Vec<T, PV>;
  pdl_trigger();
  const int token = blockIdx.x;
  const int embed = rot_dim / 2, per_head = embed / PV;
  pdl_wait();  // positions may come from advance_step_flashattn earlier in the same chain
  const T* cache = cos_sin_cache + (positions[token] + (cache_offsets ? cache_offsets[token] : 0)) * rot_dim;
  const int nq = num_heads * per_head, total = nq + (key != nullptr ? num_kv_heads * per_head : 0);
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const bool is_q = i < nq;
    const int j = is_q ? i : i - nq;
    const int head = j / per_head, off = (j % per_head) * PV;
    T* arr = (is_q ? query + token * query_stride : key + token * key_stride) + (int64_t)head * head_size;
    const H c = *reinterpret_cast<const H*>(cache + off), s = *reinterpret_cast<const H*>(cache + embed + off);
    if constexpr (NEOX) {
      V x = *reinterpret_cast<V*>(arr + off), y = *reinterpret_cast<V*>(arr + embed + off);
      V ox, oy;
#pragma unroll
      for (int e = 0; e < VEC; ++e) {
        ox.v[e] = sub_t(mul_t(x.v[e], c.v[e]), mul_t(y.v[e], s.v[e]));
        oy.v[e] = add_t(mul_t(y.v[e], c.v[e]), mul_t(x.v[e], s.v[e]));
      }
      *reinterpret_cast<V*>(arr + off) = ox;
      *reinterpret_cast<V*>(arr + embed + off) = oy;
    } else {
      V xy = *reinterpret_cast<V*>(arr + 2 * off), o;
#pragma unroll
      for (int e = 0; e < PV; ++e) {
        o.v[2 * e] = sub_t(mul_t(xy.v[2 * e], c.v[e]), mul_t(xy.v[2 * e + 1], s.v[e]));
        o.v[2 * e + 1] = add_t(mul_t(xy.v[2 * e + 1], c.v[e]), mul_t(xy.v[2 * e], s.v[e]));
      }
      *reinterpret_cast<V*>(arr + 2 * off) = o;
    }
  }
}

// Multi-step decode bookkeeping (prepare_inputs/advance_step.cu:12-66): after sampling, move every running query one
// token forward without a host round trip. One thread per query; block 0 also clears the CUDA-graph padding rows.
__global__ void __launch_bounds__(256)
    advance_step_flashattn_kernel(int num_seqs, int num_queries, int block_size, int64_t* __restrict__ input_tokens,
                                  const int64_t* __restrict__ sampled_token_ids, int64_t* __restrict__ input_positions,
                                  int* __restrict__ seq_lens, int64_t* __restrict__ slot_mapping,
                                  const int* __restrict__ block_tables, int64_t block_tables_stride) {
  pdl_trigger();
  pdl_wait();
  if (blockIdx.x == 0) {
    for (int i = num_queries + threadIdx.x; i < num_seqs; i += blockDim.x) {
      input_tokens[i] = 0;
      input_positions[i] = 0;
      slot_mapping[i] = -1;
    }
  }
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= num_queries) return;
  input_tokens[q] = sampled_token_ids[q];
  const int next_len = seq_lens[q] + 1, pos = next_len - 1;
  seq_lens[q] = next_len;
  input_positions[q] = pos;
  slot_mapping[q] = (int64_t)(block_tables[block_tables_stride * q + pos / block_size] * block_size + pos % block_size);
}

// The FlashInfer flavour of the same bookkeeping (prepare_inputs/advance_step.cu:91-167): besides the per-query
// updates it rebuilds FlashInfer's CSR view of the block tables (paged_kv_indptr / paged_kv_indices /
// paged_kv_last_page_len). The reference runs three launches with a quadratic prefix sum; here one CTA does the
// per-query pass and a chunked block scan of the page counts (kernel 1), then a grid copies each row's live pages
// into the flat list (kernel 2).
__global__ void __launch_bounds__(1024)
    advance_step_flashinfer_kernel(int num_seqs, int num_queries, int block_size, int64_t* __restrict__ input_tokens,
                                   const int64_t* __restrict__ sampled_token_ids, int64_t* __restrict__ input_positions,
                                   int* __restrict__ seq_lens, int64_t* __restrict__ slot_mapping,
                                   const int* __restrict__ block_tables, int64_t block_tables_stride,
                                   int* __restrict__ paged_kv_indptr, int* __restrict__ paged_kv_last_page_len,
                                   int* __restrict__ block_table_bound) {
  __shared__ int warp_sums[32];
  __shared__ int carry_s;
  pdl_trigger();
  pdl_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = num_queries + tid; i < num_seqs; i += blockDim.x) {
    input_tokens[i] = 0;
    input_positions[i] = 0;
    slot_mapping[i] = -1;
  }
  if (tid == 0) {
    carry_s = 0;
    paged_kv_indptr[0] = 0;
  }
  __syncthreads();
  for (int base = 0; base < num_queries; base += blockDim.x) {
    const int q = base + tid;
    int bound = 0;
    if (q < num_queries) {
      input_tokens[q] = sampled_token_ids[q];
      const int next_len = seq_lens[q] + 1, pos = next_len - 1;
      const int blk = pos / block_size, off = pos % block_size;
      seq_lens[q] = next_len;
      input_positions[q] = pos;
      paged_kv_last_page_len[q] = off + 1;
      slot_mapping[q] = (int64_t)(block_tables[block_tables_stride * q + blk] * block_size + off);
      bound = (next_len + block_size - 1) / block_size;
      block_table_bound[q] = bound;
    }
    int incl = bound;  // inclusive scan of this chunk: warp shuffles, then the warp totals
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int w = warp_sums[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, w, d);
        if (lane >= d) w += up;
      }
      warp_sums[lane] = w;
    }
    __syncthreads();
    const int carry = carry_s;
    incl += carry + (warp ? warp_sums[warp - 1] : 0);
    if (q < num_queries) paged_kv_indptr[q + 1] = incl;
    __syncthreads();
    if (tid == blockDim.x - 1) carry_s = incl;
    __syncthreads();
  }
  const int total = carry_s;  // CUDA-graph padding rows point at the end of the list
  for (int i = num_queries + 1 + tid; i <= num_seqs; i += blockDim.x) paged_kv_indptr[i] = total;
}

// One CTA per (row, 1024-column chunk): paged_kv_indices[indptr[q] + col] = block_tables[q][col] for col < bound[q].
__global__ void __launch_bounds__(256)
    advance_step_flashinfer_indices_kernel(const int* __restrict__ block_tables, int64_t block_tables_stride,
                                           int* __restrict__ paged_kv_indices, int64_t indices_capacity,
                                           const int* __restrict__ paged_kv_indptr, const int* __restrict__ block_table_bound) {
  pdl_trigger();
  pdl_wait();
  const int q = blockIdx.y;
  const int bound = block_table_bound[q];
  const int64_t start = paged_kv_indptr[q];
  const int* row = block_tables + block_tables_stride * q;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int col = (blockIdx.x * 4 + r) * 256 + threadIdx.x;
    if (col < bound && col < block_tables_stride && start + col < indices_capacity) paged_kv_indices[start + col] = row[col];
  }
}

template <typename T, bool FUSED>
int launch_norm(T* out, T* input, T* residual, const T* weight, float eps, int64_t tokens, int hidden, int64_t in_stride,
                cudaStream_t stream) {
  const bool vec_ok = hidden % (16 / (int)sizeof(T)) == 0 && aligned16(input) && aligned16(weight) &&
                      (in_stride * sizeof(T)) % 16 == 0 && (FUSED ? aligned16(residual) : aligned16(out));
  constexpr int VEC = 16 / (int)sizeof(T);
  if (vec_ok) {
    const int nvec = hidden / VEC;
    // few tokens: one vector per thread (latency); many tokens: two per thread in CTAs of at most 256 threads, five of
    // them per SM (occupancy, bytes in flight)
    if (tokens < 512 || nvec > 512) {
      const int threads = std::min(1024, std::max(32, (nvec + 31) / 32 * 32));
      APH_LAUNCH_PDL((rms_norm_kernel<T, VEC, 2, FUSED, 1024>), dim3((unsigned)tokens), dim3(threads), 0, stream, out, input, residual,
                     weight, eps, hidden, in_stride);
    } else {
      const int threads = std::max(32, ((nvec + 1) / 2 + 31) / 32 * 32);
      APH_LAUNCH_PDL((rms_norm_kernel<T, VEC, 2, FUSED, 256>), dim3((unsigned)tokens), dim3(threads), 0, stream, out, input, residual,
                     weight, eps, hidden, in_stride);
    }
  } else {
    const int threads = std::min(1024, std::max(32, (hidden + 31) / 32 * 32));
    APH_LAUNCH_PDL((rms_norm_kernel<T, 1, 1, FUSED, 1024>), dim3((unsigned)tokens), dim3(threads), 0, stream, out, input, residual,
                   weight, eps, hidden, in_stride);
  }
  return APH_OK;
}

template <typename T, int ACT, bool ACT_FIRST>
int launch_gated(T* out, const T* input, int64_t tokens, int d, float param, cudaStream_t stream) {
  constexpr int VEC = 16 / (int)sizeof(T);
  if (d % VEC == 0 && aligned16(out) && aligned16(input)) {
    APH_LAUNCH_PDL((act_and_mul_kernel<T, VEC, ACT, ACT_FIRST>), dim3((unsigned)((d / VEC + 255) / 256), (unsigned)tokens), dim3(256), 0,
                   stream, out, input, d, param);
  } else {
    APH_LAUNCH_PDL((act_and_mul_kernel<T, 1, ACT, ACT_FIRST>), dim3((unsigned)((d + 255) / 256), (unsigned)tokens), dim3(256), 0, stream,
                   out, input, d, param);
  }
  return APH_OK;
}

template <typename T>
int launch_gated_any(int act, T* out, const T* input, int64_t tokens, int d, float param, cudaStream_t stream) {
  switch (act) {
    case APH_ACT_SILU_AND_MUL: return launch_gated<T, ACT_SILU, true>(out, input, tokens, d, param, stream);
    case APH_ACT_MUL_AND_SILU: return launch_gated<T, ACT_SILU, false>(out, input, tokens, d, param, stream);
    case APH_ACT_GELU_AND_MUL: return launch_gated<T, ACT_GELU, true>(out, input, tokens, d, param, stream);
    case APH_ACT_GELU_TANH_AND_MUL: return launch_gated<T, ACT_GELU_TANH, true>(out, input, tokens, d, param, stream);
    case APH_ACT_FATRELU_AND_MUL: return launch_gated<T, ACT_FATRELU, true>(out, input, tokens, d, param, stream);
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "act_and_mul: unknown activation %d", act);
  }
}

template <typename T, int ACT>
int launch_act(T* out, const T* input, int64_t n, cudaStream_t stream) {
  constexpr int VEC = 16 / (int)sizeof(T);
  if (n % VEC == 0 && aligned16(out) && aligned16(input)) {
    APH_LAUNCH_PDL((activation_kernel<T, VEC, ACT>), dim3((unsigned)((n / VEC + 255) / 256)), dim3(256), 0, stream, out, input, n);
  } else {
    APH_LAUNCH_PDL((activation_kernel<T, 1, ACT>), dim3((unsigned)((n + 255) / 256)), dim3(256), 0, stream, out, input, n);
  }
  return APH_OK;
}

template <typename T>
int launch_act_any(int act, T* out, const T* input, int64_t n, cudaStream_t stream) {
  switch (act) {
    case APH_ACT_GELU_NEW: return launch_act<T, ACT_GELU_NEW>(out, input, n, stream);
    case APH_ACT_GELU_FAST: return launch_act<T, ACT_GELU_FAST>(out, input, n, stream);
    case APH_ACT_GELU_QUICK: return launch_act<T, ACT_GELU_QUICK>(out, input, n, stream);
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "activation: unknown activation %d", act);
  }
}

template <typename T>
int launch_rope(const int64_t* positions, T* q, T* k, const T* cache, const int64_t* offsets, int rot_dim, int64_t qs, int64_t ks,
                int nh, int nkv, int head_size, bool neox, int64_t tokens, cudaStream_t stream) {
  constexpr int VEC = 16 / (int)sizeof(T);
  const int pv = neox ? VEC : VEC / 2;
  const bool vec_ok = (rot_dim / 2) % VEC == 0 && head_size % VEC == 0 && qs % VEC == 0 && (k == nullptr || ks % VEC == 0) &&
                      aligned16(q) && (k == nullptr || aligned16(k)) && aligned16(cache);
  const int work = (nh + (k ? nkv : 0)) * (rot_dim / 2) / (vec_ok ? pv : 1);
  const int threads = std::min(512, std::max(32, (work + 31) / 32 * 32));
#define APH_ROPE(KERN, NEOX)                                                                                                  \
  APH_LAUNCH_PDL(KERN<T, NEOX>, dim3((unsigned)tokens), dim3(threads), 0, stream, positions, q, k, cache, offsets, rot_dim, qs, ks, \
                 nh, nkv, head_size)
  if (vec_ok) {
    if (neox) APH_ROPE(rotary_embedding_vec_kernel, true);
    else APH_ROPE(rotary_embedding_vec_kernel, false);
  } else {
    if (neox) APH_ROPE(rotary_embedding_kernel, true);
    else APH_ROPE(rotary_embedding_kernel, false);
  }
#undef APH_ROPE
  return APH_OK;
}

}  // namespace
}  // namespace aphb200

using namespace aphb200;

#define APH_DISPATCH_DT(dt, op, ...)                                                    \
  switch (dt) {                                                                         \
    case APH_DT_F32: { using T = float; return __VA_ARGS__; }                           \
    case APH_DT_F16: { using T = __half; return __VA_ARGS__; }                          \
    case APH_DT_BF16: { using T = __nv_bfloat16; return __VA_ARGS__; }                  \
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "%s: unknown dtype %d", op, dt); \
  }

extern "C" int aph_rms_norm(void* out, const void* input, const void* weight, float epsilon, int64_t num_tokens,
                            int64_t hidden_size, int64_t input_stride, int dtype, void* stream) {
  const char* op = "rms_norm";
  APH_CHECK(num_tokens >= 0 && hidden_size > 0 && hidden_size < (1LL << 31), APH_ERR_INVALID_ARGUMENT, "%s: bad shape", op);
  if (num_tokens == 0) return APH_OK;
  APH_CHECK(out && input && weight, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  APH_DISPATCH_DT(dtype, op, launch_norm<T, false>((T*)out, (T*)const_cast<void*>(input), nullptr, (const T*)weight, epsilon, num_tokens,
                                                   (int)hidden_size, input_stride, (cudaStream_t)stream));
}

extern "C" int aph_fused_add_rms_norm(void* input, void* residual, const void* weight, float epsilon, int64_t num_tokens,
                                      int64_t hidden_size, int dtype, void* stream) {
  const char* op = "fused_add_rms_norm";
  APH_CHECK(num_tokens >= 0 && hidden_size > 0 && hidden_size < (1LL << 31), APH_ERR_INVALID_ARGUMENT, "%s: bad shape", op);
  if (num_tokens == 0) return APH_OK;
  APH_CHECK(input && residual && weight, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  APH_DISPATCH_DT(dtype, op, launch_norm<T, true>(nullptr, (T*)input, (T*)residual, (const T*)weight, epsilon, num_tokens,
                                                  (int)hidden_size, hidden_size, (cudaStream_t)stream));
}

extern "C" int aph_act_and_mul(int act, void* out, const void* input, int64_t num_tokens, int64_t d, float param, int dtype,
                               void* stream) {
  const char* op = "act_and_mul";
  APH_CHECK(num_tokens >= 0 && d > 0 && d < (1LL << 30), APH_ERR_INVALID_ARGUMENT, "%s: bad shape", op);
  APH_CHECK(num_tokens <= 65535, APH_ERR_UNSUPPORTED, "%s: more than 65535 tokens per call", op);
  if (num_tokens == 0) return APH_OK;
  APH_CHECK(out && input, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  APH_DISPATCH_DT(dtype, op, launch_gated_any<T>(act, (T*)out, (const T*)input, num_tokens, (int)d, param, (cudaStream_t)stream));
}

extern "C" int aph_silu_and_mul(void* out, const void* input, int64_t num_tokens, int64_t d, int dtype, void* stream) {
  return aph_act_and_mul(APH_ACT_SILU_AND_MUL, out, input, num_tokens, d, 0.f, dtype, stream);
}

extern "C" int aph_activation(int act, void* out, const void* input, int64_t numel, int dtype, void* stream) {
  const char* op = "activation";
  APH_CHECK(numel >= 0 && numel < (1LL << 38), APH_ERR_INVALID_ARGUMENT, "%s: bad shape", op);
  if (numel == 0) return APH_OK;
  APH_CHECK(out && input, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  APH_DISPATCH_DT(dtype, op, launch_act_any<T>(act, (T*)out, (const T*)input, numel, (cudaStream_t)stream));
}

extern "C" int aph_rotary_embedding(const int64_t* positions, void* query, void* key, const void* cos_sin_cache,
                                    const int64_t* cos_sin_cache_offsets, int64_t num_tokens, int rot_dim, int64_t query_stride, int64_t key_stride, int num_heads,
                                    int num_kv_heads, int head_size, int is_neox, int dtype, void* stream) {
  const char* op = "rotary_embedding";
  APH_CHECK(num_tokens >= 0 && rot_dim > 0 && rot_dim % 2 == 0 && rot_dim <= head_size, APH_ERR_INVALID_ARGUMENT,
            "%s: rot_dim %d must be even and at most head_size %d", op, rot_dim, head_size);
  APH_CHECK(num_heads > 0 && num_kv_heads > 0 && num_heads % num_kv_heads == 0, APH_ERR_INVALID_ARGUMENT,
            "%s: num_heads %d is not a multiple of num_kv_heads %d", op, num_heads, num_kv_heads);
  if (num_tokens == 0) return APH_OK;
  APH_CHECK(positions && query && cos_sin_cache, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  APH_DISPATCH_DT(dtype, op, launch_rope<T>(positions, (T*)query, (T*)key, (const T*)cos_sin_cache, cos_sin_cache_offsets, rot_dim, query_stride, key_stride,
                                            num_heads, num_kv_heads, head_size, is_neox != 0, num_tokens, (cudaStream_t)stream));
}

extern "C" int aph_advance_step_flashinfer(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                                           const int64_t* sampled_token_ids, int64_t* input_positions, int32_t* seq_lens,
                                           int64_t* slot_mapping, const int32_t* block_tables, int64_t block_tables_stride,
                                           int32_t* paged_kv_indices, int64_t paged_kv_indices_len, int32_t* paged_kv_indptr,
                                           int32_t* paged_kv_last_page_len, int32_t* block_table_bound, void* stream) {
  const char* op = "advance_step_flashinfer";
  APH_CHECK(num_seqs >= 0 && num_queries >= 0 && num_queries <= num_seqs && block_size > 0, APH_ERR_INVALID_ARGUMENT,
            "%s: need 0 <= num_queries <= num_seqs and block_size > 0", op);
  APH_CHECK(block_tables_stride >= 0 && paged_kv_indices_len >= 0, APH_ERR_INVALID_ARGUMENT, "%s: negative stride or length", op);
  APH_CHECK(paged_kv_indptr, APH_ERR_INVALID_ARGUMENT, "%s: null paged_kv_indptr", op);
  APH_CHECK(num_seqs == 0 || (input_tokens && (sampled_token_ids || num_queries == 0) && input_positions && seq_lens && slot_mapping &&
                              block_tables && paged_kv_last_page_len && block_table_bound && (paged_kv_indices || paged_kv_indices_len == 0)),
            APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  APH_CHECK(num_queries <= 65535, APH_ERR_UNSUPPORTED, "%s: num_queries = %d is above the 65535 rows one launch covers", op, num_queries);
  APH_LAUNCH_PDL(advance_step_flashinfer_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream, num_seqs, num_queries, block_size,
                 input_tokens, sampled_token_ids, input_positions, seq_lens, slot_mapping, block_tables, block_tables_stride,
                 paged_kv_indptr, paged_kv_last_page_len, block_table_bound);
  if (num_queries == 0 || block_tables_stride == 0) return APH_OK;
  const unsigned chunks = (unsigned)((block_tables_stride + 1023) / 1024);
  APH_LAUNCH_PDL(advance_step_flashinfer_indices_kernel, dim3(chunks, (unsigned)num_queries), dim3(256), 0, (cudaStream_t)stream,
                 block_tables, block_tables_stride, paged_kv_indices, paged_kv_indices_len, (const int*)paged_kv_indptr,
                 (const int*)block_table_bound);
  return APH_OK;
}

extern "C" int aph_advance_step_flashattn(int num_seqs, int num_queries, int block_size, int64_t* input_tokens,
                                          const int64_t* sampled_token_ids, int64_t* input_positions, int32_t* seq_lens,
                                          int64_t* slot_mapping, const int32_t* block_tables, int64_t block_tables_stride,
                                          void* stream) {
  const char* op = "advance_step_flashattn";
  APH_CHECK(num_seqs >= 0 && num_queries >= 0 && num_queries <= num_seqs && block_size > 0, APH_ERR_INVALID_ARGUMENT,
            "%s: need 0 <= num_queries <= num_seqs and block_size > 0", op);
  if (num_seqs == 0) return APH_OK;
  APH_CHECK(input_tokens && (sampled_token_ids || num_queries == 0) && input_positions && seq_lens && slot_mapping && block_tables,
            APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  const unsigned grid = (unsigned)std::max(1, (num_queries + 255) / 256);
  APH_LAUNCH_PDL(advance_step_flashattn_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream, num_seqs, num_queries, block_size,
                 input_tokens, sampled_token_ids, input_positions, seq_lens, slot_mapping, block_tables, block_tables_stride);
  return APH_OK;
}
