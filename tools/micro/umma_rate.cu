// Microbenchmark: issue/retire rate of tcgen05.mma kind::f16 (K = 16) from one elected thread, as a function of
// the shape (M = 128, N = 16..256), the A source (tensor memory vs shared memory) and whether consecutive MMAs
// accumulate into the same D tile. Operand contents are garbage; only time matters.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../aphrodite-fork_b200/csrc/ptx.cuh"
using namespace aphb200::ptx;

// mode 0: A from TMEM, same D; 1: A from TMEM, alternating D tiles; 2: A from smem, same D
__global__ void __launch_bounds__(128) rate_kernel(int n, int mode, int iters, long long* out) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  __shared__ uint64_t bars[2];
  __shared__ uint32_t slot;
  const int w = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bars[0]), 1);
    mbar_init(smem_u32(&bars[1]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (w == 0) tmem_alloc(smem_u32(&slot), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (w == 0 || (mode == 3 && w == 1)) {
    const uint32_t idesc = umma_idesc_f16(1, 128, n);
    const uint32_t hi = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);
    const uint32_t lo = ((base >> 4) & 0x3fffu) | (1u << 16);
    const uint32_t lo_a = (((base + 65536) >> 4) & 0x3fffu) | (1u << 16);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        const uint32_t d = tm + ((mode == 1 && (kk & 1)) ? 256 : 0) + (w == 1 ? 128 : 0);
        if (mode == 2) {
          asm volatile(
              "{\n.reg .pred q;\n.reg .b64 ad, bd;\nelect.sync _|q, 0xffffffff;\nmov.b64 ad, {%1, %3};\nmov.b64 bd, {%2, %3};\n"
              "@q tcgen05.mma.cta_group::1.kind::f16 [%0], ad, bd, %4, 1;\n}\n" ::"r"(d),
              "r"(lo_a + (kk & 3) * 2), "r"(lo + (kk & 3) * 2), "r"(hi), "r"(idesc)
              : "memory");
        } else {
          tc_mma_ts_elect(d, tm + 448 + (kk & 3) * 8, lo + (kk & 3) * 2, hi, idesc, 1u);
        }
      }
    }
    long long t1 = clock64();
    tc_commit_elect(smem_u32(&bars[w]));
    mbar_wait(smem_u32(&bars[w]), 0);
    long long t2 = clock64();
    if ((threadIdx.x & 31) == 0 && w == 0 && blockIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
  }
  __syncthreads();
  if (w == 0) tmem_dealloc(tm, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int iters = 200;
  for (int mode = 0; mode < 4; ++mode)
    for (int n : {16, 64, 128}) {
      for (int grid : {148}) {
        rate_kernel<<<grid, 128, 180 * 1024>>>(n, mode, iters, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[2];
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("mode=%d (%s) N=%3d grid=%3d: issue %.1f cycles/MMA, retire %.1f cycles/MMA %s\n", mode,
               mode == 0 ? "A=TMEM same D" : mode == 1 ? "A=TMEM alt D " : mode == 2 ? "A=SMEM same D" : "2 issuers TS  ", n, grid, (double)h[0] / (iters * 8),
               (double)h[1] / (iters * 8), e == cudaSuccess ? "" : cudaGetErrorString(e));
      }
    }
  return 0;
}
