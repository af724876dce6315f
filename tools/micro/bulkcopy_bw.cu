// Microbenchmark: HBM -> shared streaming with cp.async.bulk (1-D) as a function of copy size,
// copies in flight and CTAs per SM. One warp per CTA: lane 0 issues, the warp waits on mbarriers.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../aphrodite-fork_b200/csrc/ptx.cuh"
using namespace aphb200::ptx;

template <int WARPS>
__global__ void stream_kernel(const uint8_t* src, size_t bytes_per_cta, int copy_bytes, int depth, int copies_per_stage,
                              unsigned long long* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stage_bytes = copy_bytes * copies_per_stage;
  uint8_t* my = smem + (size_t)w * depth * stage_bytes;
  const uint32_t base = smem_u32(my);
  __shared__ uint64_t bars[WARPS * 32];
  const uint32_t bar = smem_u32(&bars[w * 32]);
  if (lane == 0) for (int i = 0; i < depth; ++i) mbar_init(bar + i * 8, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();
  const size_t per_warp = bytes_per_cta / WARPS;
  const uint8_t* p = src + (size_t)blockIdx.x * bytes_per_cta + (size_t)w * per_warp;
  const int nstages = (int)(per_warp / stage_bytes);
  auto issue = [&](int i) {
    if (lane == 0) {
      const uint32_t b = bar + (i % depth) * 8;
      mbar_expect_tx(b, stage_bytes);
      for (int c = 0; c < copies_per_stage; ++c)
        bulk_g2s(base + (i % depth) * stage_bytes + c * copy_bytes, p + (size_t)i * stage_bytes + (size_t)c * copy_bytes, copy_bytes, b);
    }
  };
  for (int i = 0; i < depth && i < nstages; ++i) issue(i);
  unsigned long long acc = 0;
  for (int i = 0; i < nstages; ++i) {
    mbar_wait(bar + (i % depth) * 8, (i / depth) & 1);
    acc += my[(i % depth) * stage_bytes + lane * 4];
    __syncwarp();
    if (i + depth < nstages) issue(i + depth);
  }
  if (acc == 0x123456789ULL) *sink = acc;
}

int main() {
  const size_t total = 1ull << 30;
  uint8_t* d; unsigned long long* sink;
  cudaMalloc(&d, total); cudaMemset(d, 1, total); cudaMalloc(&sink, 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto run = [&](int warps, int ctas_per_sm, int copy_bytes, int copies_per_stage, int depth) {
    const int grid = 148 * ctas_per_sm;
    size_t per_cta = (total / grid) / (copy_bytes * copies_per_stage * warps) * (copy_bytes * copies_per_stage * warps);
    const int smem = warps * depth * copy_bytes * copies_per_stage;
    if (smem > 200 * 1024 / ctas_per_sm) return;
    float best = 1e9;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      if (warps == 1) { cudaFuncSetAttribute(stream_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        stream_kernel<1><<<grid, 32, smem>>>(d, per_cta, copy_bytes, depth, copies_per_stage, sink); }
      else { cudaFuncSetAttribute(stream_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        stream_kernel<4><<<grid, 128, smem>>>(d, per_cta, copy_bytes, depth, copies_per_stage, sink); }
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    printf("warps %d ctas/sm %d copy %6d B x%d/stage depth %2d smem/cta %6d : %7.1f GB/s  (%.1f M copies/s/SM) %s\n", warps, ctas_per_sm,
           copy_bytes, copies_per_stage, depth, smem, per_cta * grid / best / 1e6,
           per_cta * grid / (double)copy_bytes / best / 1e3 / 148, err == cudaSuccess ? "" : cudaGetErrorString(err));
  };
  for (int warps : {1, 4})
    for (int cps : {1, 2, 4})
      for (int cb : {256, 512, 1024, 2048, 4096, 8192, 16384})
        for (int cpst : {1, 8})
          for (int depth : {4, 8}) run(warps, cps, cb, cpst, depth);
  return 0;
}
