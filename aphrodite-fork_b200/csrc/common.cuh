// Shared helpers for the sm_100a kernels behind include/aphrodite_b200.h.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/aphrodite_b200.h"

namespace aphb200 {

// Thread-local error text behind aph_b200_last_error().
char* error_buffer();
int set_error(int code, const char* fmt, ...);

#define APH_CHECK(cond, code, ...)                        \
  do {                                                    \
    if (!(cond)) return ::aphb200::set_error(code, __VA_ARGS__); \
  } while (0)

#define APH_LAUNCH_CHECK()                                                               \
  do {                                                                                   \
    cudaError_t e__ = cudaGetLastError();                                                \
    if (e__ != cudaSuccess)                                                              \
      return ::aphb200::set_error(APH_ERR_CUDA, "CUDA launch failed: %s", cudaGetErrorString(e__)); \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Number of SMs of the current device (cached per device; cudaDeviceGetAttribute is not
// free and the reference pays it on every GEMM call, gptq_marlin.cu:2104-2110).
int sm_count();

// Programmatic dependent launch (PDL). Every kernel of this library is launched with programmatic stream
// serialization and (a) issues griddepcontrol.launch_dependents early, (b) issues griddepcontrol.wait before it reads
// anything an earlier kernel may have written or writes anything an earlier kernel may still read. Back-to-back
// kernels of a decode step then overlap launch latency and prologue with the previous kernel's tail; next to foreign
// kernels (torch, NCCL) the attribute degrades to ordinary stream order. APH_PDL=0 switches the attribute off.
bool pdl_enabled();
// APH_DEV_ENV=1 at process start: the development overrides (APH_GEMM_NT / _GRID, APH_PA_WAVES / _MIN_PART) are
// re-read on every call (sweep tools); otherwise they are read once and the call path never touches getenv.
bool dev_env_live();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// The same launch with thread-block clusters of `cluster_z` CTAs along z (gridDim.z must be a multiple of it).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_cluster(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster_z,
                                      Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster_z > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = 1;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = (unsigned)cluster_z;
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

#define APH_LAUNCH_PDL(...)                                                                               \
  do {                                                                                                    \
    cudaError_t e__ = ::aphb200::launch_pdl(__VA_ARGS__);                                                 \
    if (e__ != cudaSuccess)                                                                               \
      return ::aphb200::set_error(APH_ERR_CUDA, "CUDA launch failed: %s", cudaGetErrorString(e__));       \
  } while (0)

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename T>
struct ScalarTraits;
template <>
struct ScalarTraits<float> {
  static __device__ __forceinline__ float to_float(float v) { return v; }
  static __device__ __forceinline__ float from_float(float v) { return v; }
};
template <>
struct ScalarTraits<__half> {
  static __device__ __forceinline__ float to_float(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_float(float v) { return __float2half_rn(v); }
};
template <>
struct ScalarTraits<__nv_bfloat16> {
  static __device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_float(float v) { return __float2bfloat16_rn(v); }
};

// FP8 write codec, identical to the reference's
// __nv_cvt_float_to_fp8(float(x) / scale, __NV_SATFINITE, kind)
// (kernels/quantization/fp8/nvidia/quant_utils.cuh:457-489): true IEEE division, RN-even,
// saturate-to-finite. Never compile this file with --use_fast_math.
template <int KV>
__device__ __forceinline__ uint8_t fp8_encode(float x, float scale) {
  return (uint8_t)__nv_cvt_float_to_fp8(__fdiv_rn(x, scale), __NV_SATFINITE,
                                        KV == APH_KV_FP8_E5M2 ? __NV_E5M2 : __NV_E4M3);
}
// FP8 read codec: fp8 -> half (exact) -> float (quant_utils.cuh:295-300,348-359 before scaling).
template <int KV>
__device__ __forceinline__ float fp8_decode(uint8_t b) {
  __half_raw r = __nv_cvt_fp8_to_halfraw((__nv_fp8_storage_t)b, KV == APH_KV_FP8_E5M2 ? __NV_E5M2 : __NV_E4M3);
  return __half2float(__half(r));
}

}  // namespace aphb200
