// One-shot all-reduce over NVLink peer memory for the row-parallel GEMM outputs of a decode step (the
// collective half; replaces cross_device_reduce_1stage of /root/reference/kernels/all_reduce/custom_all_reduce.cuh:296-318
// and its barriers :140-200). Messages are M x hidden x 2 B (16 KiB ... 4 MiB): latency-bound, so ONE kernel does
//   (A) copy the local partial into this rank's symmetric buffer            (16-byte vectors)
//   (B) barrier with the same CTA of every peer                             (flags in peer memory, release/acquire.sys)
//   (C) read that CTA's slice from every rank's buffer over NVLink, add in fp32 in RANK ORDER, round once, store
//   (D) barrier again, so nobody overwrites a buffer a peer still reads
// Every rank adds the same values in the same order: the result is bit-identical on all ranks and from run to run.
// The symmetric buffers come from the host (torch symmetric memory / CUDA IPC): [ 4 KiB of flags | data ].
// A CTA only ever waits for the CTA with the same index on the peers, and grids are at most 64 CTAs of 512 threads, so
// all waiting CTAs are resident (no deadlock on a 148-SM part). Flags are per-CTA counters that only grow; the two
// barriers use two arrays so a fast peer's next arrival cannot be mistaken for this one (same reasoning as
// custom_all_reduce.cuh:47-60).
#include "common.cuh"

namespace aphb200 {
namespace {

constexpr int AR_MAX_BLOCKS = 64;
constexpr int AR_MAX_RANKS = 8;
constexpr int AR_SIGNAL_BYTES = 4096;

struct ArSignals {
  uint32_t start[AR_MAX_BLOCKS][AR_MAX_RANKS];
  uint32_t end[AR_MAX_BLOCKS][AR_MAX_RANKS];
};
static_assert(sizeof(ArSignals) == AR_SIGNAL_BYTES, "flag area is 4 KiB");

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

template <int WORLD>
__device__ __forceinline__ void peer_barrier(uint8_t* const* bufs, int rank, uint32_t flag, bool end) {
  __syncthreads();
  if (threadIdx.x < WORLD) {
    ArSignals* theirs = reinterpret_cast<ArSignals*>(bufs[threadIdx.x]);
    ArSignals* mine = reinterpret_cast<ArSignals*>(bufs[rank]);
    st_release_sys(end ? &theirs->end[blockIdx.x][rank] : &theirs->start[blockIdx.x][rank], flag);
    const uint32_t* w = end ? &mine->end[blockIdx.x][threadIdx.x] : &mine->start[blockIdx.x][threadIdx.x];
    // a peer that never arrives (crashed rank, mismatched call sequence) must end as a launch failure, not as a hung GPU:
    // back off after the first microseconds and trap after about a minute
    for (uint32_t spins = 0; ld_acquire_sys(w) != flag; ++spins) {
      if (spins > 4096) __nanosleep(256);
      if (spins > (1u << 28)) __trap();
    }
  }
  __syncthreads();
}

template <typename T, int WORLD>
__global__ void __launch_bounds__(512, 1)
    one_shot_all_reduce_kernel(T* out, const T* in, uint8_t* const* __restrict__ peer_bufs,  // out may alias in
                               uint32_t* __restrict__ counters, int rank, int64_t n_vec) {
  constexpr int VEC = 16 / (int)sizeof(T);
  using V = uint4;
  __shared__ uint8_t* s_bufs[WORLD];
  pdl_trigger();
  if (threadIdx.x < WORLD) s_bufs[threadIdx.x] = peer_bufs[threadIdx.x];  // immutable table
  const uint32_t flag = counters[blockIdx.x] + 1;                          // only this CTA ever writes its counter
  __syncthreads();
  pdl_wait();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x, first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  V* own = reinterpret_cast<V*>(s_bufs[rank] + AR_SIGNAL_BYTES);
  for (int64_t i = first; i < n_vec; i += stride) own[i] = reinterpret_cast<const V*>(in)[i];
  peer_barrier<WORLD>(s_bufs, rank, flag, false);
  for (int64_t i = first; i < n_vec; i += stride) {
    V v[WORLD];
#pragma unroll
    for (int r = 0; r < WORLD; ++r) v[r] = reinterpret_cast<const V*>(s_bufs[r] + AR_SIGNAL_BYTES)[i];  // all loads in flight
    float acc[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] = ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&v[0])[e]);
#pragma unroll
    for (int r = 1; r < WORLD; ++r)
#pragma unroll
      for (int e = 0; e < VEC; ++e) acc[e] += ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&v[r])[e]);
    V o;
#pragma unroll
    for (int e = 0; e < VEC; ++e) reinterpret_cast<T*>(&o)[e] = ScalarTraits<T>::from_float(acc[e]);
    reinterpret_cast<V*>(out)[i] = o;
  }
  peer_barrier<WORLD>(s_bufs, rank, flag, true);
  if (threadIdx.x == 0) counters[blockIdx.x] = flag;
}

template <typename T>
int launch_one_shot(void* out, const void* in, const void* const* peer_bufs, uint32_t* counters, int rank, int world, int64_t n_vec,
                    int blocks, cudaStream_t stream) {
#define APH_AR(W)                                                                                                             \
  case W:                                                                                                                     \
    APH_LAUNCH_PDL((one_shot_all_reduce_kernel<T, W>), dim3((unsigned)blocks), dim3(512), (size_t)0, stream, (T*)out, (const T*)in, \
                   (uint8_t* const*)peer_bufs, counters, rank, n_vec);                                                        \
    return APH_OK
  switch (world) {
    APH_AR(2);
    APH_AR(4);
    APH_AR(6);
    APH_AR(8);
    default: return set_error(APH_ERR_UNSUPPORTED, "one_shot_all_reduce: world size %d (2, 4, 6 or 8)", world);
  }
#undef APH_AR
}

}  // namespace
}  // namespace aphb200

using namespace aphb200;

extern "C" size_t aph_all_reduce_signal_bytes(void) { return AR_SIGNAL_BYTES; }
extern "C" int aph_all_reduce_max_blocks(void) { return AR_MAX_BLOCKS; }

extern "C" int aph_one_shot_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank,
                                       int world, int64_t numel, int dtype, int max_blocks, void* stream) {
  const char* op = "one_shot_all_reduce";
  APH_CHECK(world >= 2 && world <= AR_MAX_RANKS && rank >= 0 && rank < world, APH_ERR_INVALID_ARGUMENT, "%s: rank %d of %d", op, rank,
            world);
  APH_CHECK(numel >= 0, APH_ERR_INVALID_ARGUMENT, "%s: negative size", op);
  if (numel == 0) return APH_OK;
  APH_CHECK(out && in && peer_bufs_dev && counters, APH_ERR_INVALID_ARGUMENT, "%s: null pointer", op);
  const int esz = dtype == APH_DT_F32 ? 4 : 2;
  APH_CHECK((numel * esz) % 16 == 0 && aligned16(out) && aligned16(in), APH_ERR_INVALID_ARGUMENT,
            "%s: the message must be a multiple of 16 bytes and 16-byte aligned", op);
  const int64_t n_vec = numel * esz / 16;
  int blocks = (int)std::min<int64_t>((n_vec + 511) / 512, std::min(max_blocks > 0 ? max_blocks : AR_MAX_BLOCKS, AR_MAX_BLOCKS));
  blocks = std::max(blocks, 1);
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case APH_DT_F32: return launch_one_shot<float>(out, in, peer_bufs_dev, counters, rank, world, n_vec, blocks, s);
    case APH_DT_F16: return launch_one_shot<__half>(out, in, peer_bufs_dev, counters, rank, world, n_vec, blocks, s);
    case APH_DT_BF16: return launch_one_shot<__nv_bfloat16>(out, in, peer_bufs_dev, counters, rank, world, n_vec, blocks, s);
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "%s: unknown dtype %d", op, dtype);
  }
}
