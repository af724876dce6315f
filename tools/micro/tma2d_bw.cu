// Microbenchmark: HBM -> shared streaming of a Marlin-format weight matrix ([K/16][N] 8-byte words) with
// 2-D tiled TMA loads, as a function of the box shape (NT columns x rows = 16 KiB), ring depth, CTAs per SM
// and the k split. Producer = lane 0 of warp 1, consumer = warp 0 (waits, touches the tile, releases).
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../aphrodite-fork_b200/csrc/ptx.cuh"
using namespace aphb200::ptx;

__global__ void __launch_bounds__(64) stream2d(const __grid_constant__ CUtensorMap tm, int nt, int box_rows, int depth,
                                                 int split, int stages_total, unsigned long long* sink) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* smem = raw + (base - smem_u32(raw));
  __shared__ uint64_t bars[32];
  const uint32_t full = smem_u32(&bars[0]), empty = smem_u32(&bars[16]);
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int stage_bytes = nt * 8 * box_rows;
  if (threadIdx.x == 0) {
    for (int i = 0; i < depth; ++i) { mbar_init(full + i * 8, 1); mbar_init(empty + i * 8, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int slice = blockIdx.x / split, chunk = blockIdx.x % split;
  const int per = (stages_total + split - 1) / split;
  const int kb = chunk * per;
  const int nk = min(stages_total, kb + per) - kb;
  if (w == 1) {
    if (lane != 0) return;
    const uint64_t pol = policy_evict_first();
    for (int it = 0; it < nk; ++it) {
      const int st = it % depth;
      mbar_wait(empty + st * 8, ((it / depth) & 1) ^ 1);
      mbar_expect_tx(full + st * 8, stage_bytes);
      tma_load_2d_hint(base + st * stage_bytes, &tm, slice * nt, (kb + it) * box_rows, full + st * 8, pol);
    }
    return;
  }
  unsigned long long acc = 0;
  for (int it = 0; it < nk; ++it) {
    const int st = it % depth;
    mbar_wait(full + st * 8, (it / depth) & 1);
    acc += smem[st * stage_bytes + lane * 16];
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + st * 8);
  }
  if (acc == 0x123456789ULL) *sink = acc;
}

typedef CUresult (*EncFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                          const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                          CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q);
  EncFn enc = (EncFn)sym;
  const size_t total = 1ull << 30;
  uint8_t* d;
  unsigned long long* sink;
  cudaMalloc(&d, total);
  cudaMemset(d, 1, total);
  cudaMalloc(&sink, 8);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaFuncSetAttribute(stream2d, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  auto run = [&](int N, int K, int nt, int depth, int split, int promo) {
    const int box_rows = 16384 / (nt * 8);
    const int rows = K / 16;
    const int stages_total = rows / box_rows;
    // several matrices back to back so that one timed launch sequence reads > L2
    const size_t mat_bytes = (size_t)rows * N * 8;
    const int nmat = (int)(total / mat_bytes);
    const int reps = nmat < 8 ? nmat : 8;
    const int grid = (N / nt) * split;
    const int smem = depth * 16384 + 1024;
    float best = 1e9f;
    for (int trial = 0; trial < 3; ++trial) {
      cudaEventRecord(e0);
      for (int r = 0; r < reps; ++r) {
        CUtensorMap tm;
        const cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)rows};
        const cuuint64_t strides[1] = {(cuuint64_t)N * 8};
        const cuuint32_t box[2] = {(cuuint32_t)nt, (cuuint32_t)box_rows};
        const cuuint32_t es[2] = {1, 1};
        enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, d + (size_t)r * mat_bytes, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        stream2d<<<grid, 64, smem>>>(tm, nt, box_rows, depth, split, stages_total, sink);
      }
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    printf("N=%5d K=%5d nt=%3d box_rows=%2d depth=%d split=%d grid=%4d promo=%d : %7.2f us/launch  %7.1f GB/s %s\n", N, K, nt, box_rows,
           depth, split, grid, promo, best * 1e3 / reps, mat_bytes * reps / (best * 1e-3) / 1e9, err == cudaSuccess ? "" : cudaGetErrorString(err));
  };
  const int shapes[4][2] = {{28672, 4096}, {4096, 14336}, {6144, 4096}, {4096, 4096}};
  for (auto& sh : shapes) {
    const int N = sh[0], K = sh[1];
    for (int nt : {64, 128, 256})
      for (int depth : {2, 4, 8}) {
        const int slices = N / nt;
        for (int target : {148, 296, 592}) {
          int split = (target + slices - 1) / slices;
          if (split < 1) split = 1;
          const int stages_total = K / 16 / (16384 / (nt * 8));
          if (split > stages_total) continue;
          run(N, K, nt, depth, split, 2);
        }
      }
  }
  run(28672, 4096, 256, 4, 2, 0);
  run(28672, 4096, 256, 4, 2, 1);
  run(28672, 4096, 256, 4, 2, 3);
  return 0;
}
