// gptq_marlin_repack / awq_marlin_repack for sm_100a: checkpoint weight formats -> the Marlin wire format the GEMM
// reads. Replaces /root/reference/kernels/quantization/gptq_marlin/gptq_marlin_repack.cu:7-322
// and awq_marlin_repack.cu:7-268. Results are bit-identical with the reference kernels and with the reference's
// Python layout code (marlin_utils_test.py:30-92); the implementation is the layout as closed-form index arithmetic,
// one thread per output word, instead of the reference's staged shared-memory pipeline: it runs once per weight at
// load time and is bound by L2 gathers.
//
//   output   b [K/16][N*16/pack] int32, pack = 32 / num_bits. Row kt = k tile; the row is a sequence of blocks of 1024
//            values (16 k x 64 n), block nb covers columns [64 nb, 64 nb + 64). Inside a block the un-interleaved
//            position u = lane*32 + tile*8 + e holds W[16 kt + row][64 nb + 16 tile + col] with
//            row = 2 (lane % 4) + (e % 4) % 2 + 8 ((e % 4) / 2),  col = lane / 4 + 8 (e / 4)
//            (the values one lane of mma.m16n8k16 needs), and output value q = pack*r + j of the block is
//            u = pack*r + interleave[j], interleave = {0,2,4,6,1,3,5,7} (4 bit) / {0,2,1,3} (8 bit).
//   GPTQ in  [K/pack][N] int32, value k of a word = bits [num_bits (k % pack) ...]; optional act-order `perm`:
//            output row k reads input row perm[k] (the reference's sort_indices).
//   AWQ in   [K][N/pack] int32, value at bit position i of a word = column pack*(n/pack) + interleave[i].
#include "common.cuh"

namespace aphb200 {
namespace {

template <int BITS, bool AWQ>
__global__ void __launch_bounds__(256)
    marlin_repack_kernel(int32_t* __restrict__ out, const uint32_t* __restrict__ in, const int32_t* __restrict__ perm,
                         int size_k, int size_n) {
  constexpr int PACK = 32 / BITS;
  constexpr uint32_t MASK = (1u << BITS) - 1u;
  const int words_per_row = size_n * 16 / PACK;
  const int kt = blockIdx.y;
  for (int wd = blockIdx.x * blockDim.x + threadIdx.x; wd < words_per_row; wd += gridDim.x * blockDim.x) {
    const int nb = wd / (1024 / PACK), r = wd % (1024 / PACK);
    uint32_t word = 0;
#pragma unroll
    for (int j = 0; j < PACK; ++j) {
      const int il = BITS == 4 ? ((j & 3) * 2 + (j >> 2)) : ((j & 1) * 2 + (j >> 1));  // {0,2,4,6,1,3,5,7} / {0,2,1,3}
      const int u = PACK * r + il;
      const int lane = u >> 5, tile = (u >> 3) & 3, e = u & 7;
      const int row = 2 * (lane & 3) + (e & 1) + 8 * ((e & 3) >> 1);
      const int col = (lane >> 2) + 8 * (e >> 2);
      int k = kt * 16 + row;
      const int n = nb * 64 + tile * 16 + col;
      uint32_t v;
      if constexpr (AWQ) {
        // column m = n % PACK sits at bit position inverse_interleave[m]
        const int m = n % PACK;
        const int pos = BITS == 4 ? ((m & 1) * 4 + (m >> 1)) : ((m & 1) * 2 + (m >> 1));
        v = (in[(int64_t)k * (size_n / PACK) + n / PACK] >> (BITS * pos)) & MASK;
      } else {
        if (perm != nullptr) k = perm[k];
        v = (in[(int64_t)(k / PACK) * size_n + n] >> (BITS * (k % PACK))) & MASK;
      }
      word |= v << (BITS * j);
    }
    out[(int64_t)kt * words_per_row + wd] = (int32_t)word;
  }
}

// Act-order: out[m][k] = a[m][perm[k]] (reference permute_cols_kernel, gptq_marlin.cu:459-515). With the weight rows
// sorted by group (the repack above with `perm`) and a full K this turns an act-order GEMM into the plain grouped one.
// 2-byte elements; a thread gathers 8 and stores them as one 16-byte vector.
__global__ void __launch_bounds__(256)
    permute_cols_kernel(uint16_t* __restrict__ out, const uint16_t* __restrict__ a, const int32_t* __restrict__ perm, int size_k,
                        int64_t lda) {
  pdl_trigger();
  const int64_t row = blockIdx.y;
  const int k0 = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (k0 >= size_k) return;
  int src[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) src[j] = perm[k0 + j];  // immutable: before the wait
  pdl_wait();
  alignas(16) uint16_t v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = a[row * lda + src[j]];
  *reinterpret_cast<uint4*>(out + row * size_k + k0) = *reinterpret_cast<const uint4*>(v);
}

template <bool AWQ>
int launch_repack(int32_t* out, const int32_t* in, const int32_t* perm, int64_t size_k, int64_t size_n, int num_bits,
                  cudaStream_t stream) {
  const int pack = 32 / num_bits;
  const int64_t words_per_row = size_n * 16 / pack;
  dim3 grid((unsigned)((words_per_row + 255) / 256), (unsigned)(size_k / 16));
  if (num_bits == 4)
    marlin_repack_kernel<4, AWQ><<<grid, 256, 0, stream>>>(out, (const uint32_t*)in, perm, (int)size_k, (int)size_n);
  else
    marlin_repack_kernel<8, AWQ><<<grid, 256, 0, stream>>>(out, (const uint32_t*)in, perm, (int)size_k, (int)size_n);
  APH_LAUNCH_CHECK();
  return APH_OK;
}

int check_repack(const char* op, const void* out, const void* in, int64_t size_k, int64_t size_n, int num_bits) {
  // the reference's checks (gptq_marlin_repack.cu:247-262, awq_marlin_repack.cu:196-211)
  APH_CHECK(num_bits == 4 || num_bits == 8, APH_ERR_INVALID_ARGUMENT, "%s: num_bits must be 4 or 8. Got = %d", op, num_bits);
  APH_CHECK(size_k > 0 && size_k % 16 == 0, APH_ERR_INVALID_ARGUMENT, "%s: size_k = %lld is not divisible by tile_k_size = 16", op,
            (long long)size_k);
  APH_CHECK(size_n > 0 && size_n % 64 == 0, APH_ERR_INVALID_ARGUMENT, "%s: size_n = %lld is not divisible by tile_n_size = 64", op,
            (long long)size_n);
  APH_CHECK(size_k / 16 <= 65535 && size_k < (1LL << 31) && size_n < (1LL << 27), APH_ERR_UNSUPPORTED, "%s: matrix too large", op);
  APH_CHECK(out != nullptr && in != nullptr, APH_ERR_INVALID_ARGUMENT, "%s: null tensor", op);
  return APH_OK;
}

}  // namespace
}  // namespace aphb200

using namespace aphb200;

extern "C" int aph_gptq_marlin_repack(int32_t* out, const int32_t* b_q_weight, const int32_t* perm, int64_t size_k,
                                      int64_t size_n, int num_bits, void* stream) {
  const int st = check_repack("gptq_marlin_repack", out, b_q_weight, size_k, size_n, num_bits);
  if (st != APH_OK) return st;
  return launch_repack<false>(out, b_q_weight, perm, size_k, size_n, num_bits, (cudaStream_t)stream);
}

extern "C" int aph_awq_marlin_repack(int32_t* out, const int32_t* b_q_weight, int64_t size_k, int64_t size_n, int num_bits,
                                     void* stream) {
  const int st = check_repack("awq_marlin_repack", out, b_q_weight, size_k, size_n, num_bits);
  if (st != APH_OK) return st;
  return launch_repack<true>(out, b_q_weight, nullptr, size_k, size_n, num_bits, (cudaStream_t)stream);
}

extern "C" int aph_permute_cols(void* out, const void* a, const int32_t* perm, int64_t size_m, int64_t size_k, int64_t lda,
                                void* stream) {
  const char* op = "permute_cols";
  APH_CHECK(size_m >= 0 && size_m <= 65535 && size_k > 0 && size_k % 8 == 0 && size_k < (1LL << 31), APH_ERR_INVALID_ARGUMENT,
            "%s: size_k must be a multiple of 8 and size_m at most 65535", op);
  if (size_m == 0) return APH_OK;
  APH_CHECK(out && a && perm && aligned16(out), APH_ERR_INVALID_ARGUMENT, "%s: null or misaligned tensor", op);
  APH_LAUNCH_PDL(permute_cols_kernel, dim3((unsigned)((size_k / 8 + 255) / 256), (unsigned)size_m), dim3(256), (size_t)0,
                 (cudaStream_t)stream, (uint16_t*)out, (const uint16_t*)a, perm, (int)size_k, lda);
  return APH_OK;
}
