// Microbenchmark: issue rate of the legacy mma.sync.m16n8k16 (bf16 -> f32) on sm_100a, per SM sub-partition, as a
// function of resident warps per SM and of the number of independent accumulator chains per warp. This is the
// pipe the register-accumulator W4A16 GEMM (M <= 16) runs on; the result bounds its bytes per clock.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_sync_rate mma_sync_rate.cu && ./mma_sync_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void rate_kernel(int iters, long long* out, float* sink) {
  float acc[CHAINS][4];
#pragma unroll
  for (int c = 0; c < CHAINS; ++c) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0.f;
  uint32_t a0 = threadIdx.x, a1 = a0 * 3, a2 = a0 * 5, a3 = a0 * 7, b0 = 0x3f803f80u, b1 = 0x3f803f80u;
  __syncthreads();
  const long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int c = 0; c < CHAINS; ++c)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+f"(acc[c][0]), "+f"(acc[c][1]), "+f"(acc[c][2]), "+f"(acc[c][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < CHAINS; ++c) s += acc[c][0] + acc[c][1] + acc[c][2] + acc[c][3];
  if (s == 12345.f) sink[0] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = t1 - t0;
}

template <int CHAINS>
void run(int warps_per_sm, long long* d_out, float* sink) {
  const int iters = 4096;
  rate_kernel<CHAINS><<<148, warps_per_sm * 32>>>(iters, d_out, sink);
  rate_kernel<CHAINS><<<148, warps_per_sm * 32>>>(iters, d_out, sink);
  cudaDeviceSynchronize();
  long long h = 0;
  cudaMemcpy(&h, d_out, 8, cudaMemcpyDeviceToHost);
  const double mmas_per_smsp = (double)iters * CHAINS * warps_per_sm / 4.0;
  const double clk_per_mma = (double)h / mmas_per_smsp;
  printf("warps/SM %2d chains %d: %8lld clk, %.2f clk per MMA per sub-partition, %.0f TFLOP/s dense-equivalent at 1.965 GHz\n",
         warps_per_sm, CHAINS, h, clk_per_mma, 4096.0 / clk_per_mma * 4 * 148 * 1.965e9 / 1e12);
}

int main() {
  long long* d_out;
  float* sink;
  cudaMalloc(&d_out, 8);
  cudaMalloc(&sink, 4);
  for (int w : {4, 8, 16, 32}) {
    run<1>(w, d_out, sink);
    run<2>(w, d_out, sink);
    run<5>(w, d_out, sink);
    run<8>(w, d_out, sink);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
