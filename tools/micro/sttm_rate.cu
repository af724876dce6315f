// Microbenchmark: throughput of tcgen05.st (registers -> tensor memory) per SM as a function of the shape
// (16x128b.x8 = what the W4A16 dequantisers issue, 32x32b.x16, 16x256b.x4: 2 KiB per warp instruction each), the
// number of warps storing at once, and how many stores go out between tcgen05.wait::st. Mode bit 8 adds an MMA
// issuer that keeps reading A from the same tensor memory (8 x K=16, N=64 per loop) to show the port contention.
// One CTA; build: nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/sttm_rate tools/micro/sttm_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../aphrodite-fork_b200/csrc/ptx.cuh"
using namespace aphb200::ptx;

__device__ __forceinline__ void st_32x32b_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void st_16x256b_x4(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.16x256b.x4.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

// shape 0: 16x128b.x8 (two per 32-lane quadrant slice), 1: 32x32b.x16, 2: 16x256b.x4
__global__ void __launch_bounds__(1024) sttm_kernel(int shape, int per_wait, int iters, int with_mma, long long* out) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint32_t slot;
  __shared__ uint64_t bars[4];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nst = (blockDim.x >> 5) - 1;  // the last warp is the MMA issuer (idle without with_mma)
  if (threadIdx.x == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bars[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (w == 0) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  long long t0 = 0, t1 = 0;
  if (w < nst) {
    const int qd = w & 3;
    uint32_t r[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) r[j] = threadIdx.x * 16 + j;
    const uint32_t lane_base = (uint32_t)(32 * qd) << 16;
    __syncwarp();
    t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      for (int k = 0; k < per_wait; ++k) {
        const uint32_t col = 64 + (uint32_t)(((w >> 2) * 64 + k * 32) & 255);
        if (shape == 0) {
          tmem_st_16x128b_x8(tm + lane_base + col, r);
          tmem_st_16x128b_x8(tm + lane_base + (16u << 16) + col, r);
        } else if (shape == 1) {
          st_32x32b_x16(tm + lane_base + col, r);
          st_32x32b_x16(tm + lane_base + col + 16, r);
        } else {
          st_16x256b_x4(tm + lane_base + col, r);
          st_16x256b_x4(tm + lane_base + (16u << 16) + col, r);
        }
      }
      tmem_wait_st();
    }
    t1 = clock64();
    if (lane == 0) {
      out[2 * w] = t1 - t0;
      out[2 * w + 1] = (long long)iters * per_wait * 2 * 2048;  // bytes this warp stored
    }
  } else if (with_mma) {
    const uint32_t idesc = umma_idesc_f16(1, 128, 64);
    const uint32_t hi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);
    const uint32_t lo = ((base >> 4) & 0x3fffu) | (1u << 16);
    for (int i = 0; i < iters; ++i) {
      const int sl = i & 3;
      if (i >= 4) mbar_wait(smem_u32(&bars[sl]), ((i >> 2) - 1) & 1);  // the MMAs that used this slot have retired
      tc_mma_ts_stage8(tm, tm + 64 + sl * 64, lo, hi, idesc, 1u, (uint32_t)(8192 >> 4));
      tc_commit_elect(smem_u32(&bars[sl]));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (w == 0) tmem_dealloc(tm, 512);
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 64 * sizeof(long long));
  cudaFuncSetAttribute(sttm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const char* names[3] = {"16x128b.x8", "32x32b.x16", "16x256b.x4"};
  for (int with_mma = 0; with_mma <= 1; ++with_mma)
    for (int shape = 0; shape < 3; ++shape)
      for (int warps : {4, 8, 16})
        for (int per_wait : {1, 4}) {
          const int iters = 2000;
          for (int rep = 0; rep < 2; ++rep) {
            sttm_kernel<<<1, (warps + 1) * 32, 100 * 1024>>>(shape, per_wait, iters, with_mma, out);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) {
              printf("error %s\n", cudaGetErrorString(e));
              return 1;
            }
          }
          long long cyc = 0, bytes = 0;
          for (int w = 0; w < warps; ++w) {
            cyc = out[2 * w] > cyc ? out[2 * w] : cyc;
            bytes += out[2 * w + 1];
          }
          printf("{\"shape\": \"%s\", \"warps\": %d, \"stores_per_wait\": %d, \"mma\": %d, \"cycles\": %lld, \"bytes_per_cycle\": %.1f, "
                 "\"cycles_per_2KiB_store_per_warp\": %.1f}\n",
                 names[shape], warps, 2 * per_wait, with_mma, cyc, (double)bytes / cyc, (double)cyc / (iters * per_wait * 2));
        }
  return 0;
}
