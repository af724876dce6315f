// W4A16 GEMM for act-order (desc_act) weights whose quantisation groups do NOT follow the k order of the shard:
// the row-parallel shard of a desc_act checkpoint (is_k_full = False). Replaces the has_act_order paths of the
// reference kernel (/root/reference/kernels/quantization/gptq_marlin/gptq_marlin.cu:693, 921-985 g_idx staging,
// 1280-1300 scale4 per-row scales, 1825).
//
// Rows of the (sorted) shard carry an arbitrary, non-decreasing group index g_idx[k] into the scale table of the
// WHOLE weight, so a scale cannot be applied to a group's sum; like the reference, every weight is scaled on its own
// (w = (q - 8) * s[g_idx[k], n], rounded to the 16-bit type) before the MMA. This is the rare path - one launch,
// no pipeline, plain loads through L1/L2: the weight tile is the A operand of mma.sync.m16n8k16 as in
// marlin_gemm_kernel.cuh, a CTA owns 64 output columns x 16 batch rows, its 8 warps split the k16-rows and meet in
// shared memory. Weights that do follow the k order (is_k_full) never come here: they run the grouped kernels.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace aphb200 {
namespace mgao {

using namespace ::aphb200::ptx;
constexpr int WARPS = 8;
constexpr int MP = 16;  // batch rows per CTA

template <typename scalar_t>
__global__ void __launch_bounds__(WARPS * 32)
    marlin_actorder_kernel(scalar_t* __restrict__ c, const scalar_t* __restrict__ a, const uint4* __restrict__ b,
                           const scalar_t* __restrict__ s, const int32_t* __restrict__ g_idx, int M, int N, int K, int64_t lda) {
  constexpr bool F16 = std::is_same<scalar_t, __half>::value;
  constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;
  __shared__ float red[WARPS][MP][64 + 4];
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
  const int slice = blockIdx.x, m0 = blockIdx.y * MP;
  float acc[4][2][4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int bb = 0; bb < 2; ++bb)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[j][bb][e] = 0.f;
  auto sub2 = [](uint32_t x, uint32_t cc) -> uint32_t {
    if constexpr (F16) {
      __half2 v = __hsub2(*reinterpret_cast<__half2*>(&x), *reinterpret_cast<__half2*>(&cc));
      return *reinterpret_cast<uint32_t*>(&v);
    } else {
      __nv_bfloat162 v = __hsub2(*reinterpret_cast<__nv_bfloat162*>(&x), *reinterpret_cast<__nv_bfloat162*>(&cc));
      return *reinterpret_cast<uint32_t*>(&v);
    }
  };
  auto mul2 = [](uint32_t x, uint32_t y) -> uint32_t {
    if constexpr (F16) {
      __half2 v = __hmul2(*reinterpret_cast<__half2*>(&x), *reinterpret_cast<__half2*>(&y));
      return *reinterpret_cast<uint32_t*>(&v);
    } else {
      __nv_bfloat162 v = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&x), *reinterpret_cast<__nv_bfloat162*>(&y));
      return *reinterpret_cast<uint32_t*>(&v);
    }
  };
  const int k_rows = K / 16;
  for (int kr = w; kr < k_rows; kr += WARPS) {
    // one packed word per 16-column tile: nibble pairs (0,4) (2,6) (1,5) (3,7) are the registers a0 a1 a2 a3
    const uint4 qv = __ldg(b + ((size_t)kr * (size_t)(N / 2) + (size_t)slice * 32 + lane));
    const uint32_t q[4] = {qv.x, qv.y, qv.z, qv.w};
    // groups of this thread's four k values (2t, 2t+1, 2t+8, 2t+9 of the k16-row) and their scale rows: the 8 scales
    // (4 tiles x rows g / g+8) of one (group, thread) are 16 contiguous bytes of the Marlin-permuted table
    const int kb = kr * 16 + 2 * t;
    uint32_t sv[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int gi = __ldg(g_idx + kb + (i & 1) + (i >> 1) * 8);
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(s + (size_t)gi * N + slice * 64 + g * 8));
      sv[i][0] = v.x;
      sv[i][1] = v.y;
      sv[i][2] = v.z;
      sv[i][3] = v.w;
    }
    // activations: B fragments of two 8-row blocks
    uint32_t bf[2][2];
#pragma unroll
    for (int bb = 0; bb < 2; ++bb) {
      const int m = m0 + bb * 8 + g;
      if (m < M) {
        const scalar_t* ap = a + (size_t)m * lda + kb;
        bf[bb][0] = *reinterpret_cast<const uint32_t*>(ap);
        bf[bb][1] = *reinterpret_cast<const uint32_t*>(ap + 8);
      } else {
        bf[bb][0] = bf[bb][1] = 0u;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      // scale pairs (k even, k odd) for row g (low half of word j) and row g + 8 (high half)
      const uint32_t s_lo_g = __byte_perm(sv[0][j], sv[1][j], 0x5410), s_lo_g8 = __byte_perm(sv[0][j], sv[1][j], 0x7632);
      const uint32_t s_hi_g = __byte_perm(sv[2][j], sv[3][j], 0x5410), s_hi_g8 = __byte_perm(sv[2][j], sv[3][j], 0x7632);
      const uint32_t cz = MAGIC | 0x00080008u;
      const uint32_t a0 = mul2(sub2(and_or(q[j], 0x000f000fu, MAGIC), cz), s_lo_g);
      const uint32_t a2 = mul2(sub2(and_or(q[j] >> 4, 0x000f000fu, MAGIC), cz), s_hi_g);
      const uint32_t a1 = mul2(sub2(and_or(q[j] >> 8, 0x000f000fu, MAGIC), cz), s_lo_g8);
      const uint32_t a3 = mul2(sub2(and_or(q[j] >> 12, 0x000f000fu, MAGIC), cz), s_hi_g8);
#pragma unroll
      for (int bb = 0; bb < 2; ++bb) mma16816<F16>(acc[j][bb], a0, a1, a2, a3, bf[bb][0], bf[bb][1]);
    }
  }
  // accumulator element e of tile j, block bb: column 16 j + g (+ 8 for e >= 2), batch row 8 bb + 2 t + (e & 1)
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int bb = 0; bb < 2; ++bb)
#pragma unroll
      for (int e = 0; e < 4; ++e) red[w][bb * 8 + 2 * t + (e & 1)][16 * j + g + (e >> 1) * 8] = acc[j][bb][e];
  __syncthreads();
  for (int idx = threadIdx.x; idx < MP * 64; idx += WARPS * 32) {
    const int m = idx / 64, n = idx % 64;
    if (m0 + m >= M) continue;
    float v = 0.f;
#pragma unroll
    for (int ww = 0; ww < WARPS; ++ww) v += red[ww][m][n];
    c[(size_t)(m0 + m) * N + slice * 64 + n] = ScalarTraits<scalar_t>::from_float(v);
  }
}

template <typename scalar_t>
static int launch(void* c, const void* a, const void* b, const void* s, const int32_t* g_idx, int64_t M, int64_t N, int64_t K,
                  int64_t lda, cudaStream_t stream) {
  dim3 grid((unsigned)(N / 64), (unsigned)((M + MP - 1) / MP));
  APH_LAUNCH_PDL(marlin_actorder_kernel<scalar_t>, grid, dim3(WARPS * 32), (size_t)0, stream, (scalar_t*)c, (const scalar_t*)a,
                 (const uint4*)b, (const scalar_t*)s, g_idx, (int)M, (int)N, (int)K, lda);
  return APH_OK;
}

}  // namespace mgao
}  // namespace aphb200
