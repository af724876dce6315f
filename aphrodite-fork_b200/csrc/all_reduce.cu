// One-shot all-reduce over NVLink peer memory for the row-parallel GEMM outputs of a decode step (the
// collective half; replaces cross_device_reduce_1stage of /root/reference/kernels/all_reduce/custom_all_reduce.cuh:296-318
// and its barriers :140-200). Messages are M x hidden x 2 B (16 KiB ... 4 MiB): latency-bound, so ONE kernel does
//   (A) PUSH the local partial into slot `rank` of every rank's symmetric buffer (posted 16-byte stores over NVLink: no
//       round trip per access, unlike the reference's peer loads)
//   (B) one barrier with the same CTA of every peer                         (flags in peer memory, release/acquire.sys)
//   (C) add the `world` slots of the LOCAL buffer in fp32 in RANK ORDER, round once, store
// Every rank adds the same values in the same order: the result is bit-identical on all ranks and from run to run.
// The buffers are double: call n uses half n & 1, so a fast rank's pushes of call n + 1 never touch what a slow rank
// still reads in call n, and nobody can start call n + 2 before every rank has passed the barrier of call n + 1, i.e.
// finished call n. That removes the reference's second barrier. Flags are per-CTA counters that only grow; a waiter
// accepts any value >= its own flag because a fast peer may already have published the next one.
// Layout of a rank's buffer (from the host: torch symmetric memory / CUDA IPC): [ 4 KiB flags | 2 halves x world slots
// x slot_bytes ]. A CTA only waits for the CTA with the same index on the peers and grids are at most 128 CTAs of 512
// threads, so every waiting CTA is resident (no deadlock on a 148-SM part).
#include "common.cuh"

namespace aphb200 {
namespace {

constexpr int AR_MAX_BLOCKS = 128;
constexpr int AR_MAX_RANKS = 8;
constexpr int AR_SIGNAL_BYTES = 4096;

struct ArSignals {
  uint32_t arrived[AR_MAX_BLOCKS][AR_MAX_RANKS];
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

template <typename T, int WORLD>
__global__ void __launch_bounds__(512, 1)
    one_shot_all_reduce_kernel(T* out, const T* in, uint8_t* const* __restrict__ peer_bufs,  // out may alias in
                               uint32_t* __restrict__ counters, int rank, int64_t n_vec, int64_t slot_bytes) {
  constexpr int VEC = 16 / (int)sizeof(T);
  using V = uint4;
  __shared__ uint8_t* s_bufs[WORLD];
  pdl_trigger();
  if (threadIdx.x < WORLD) s_bufs[threadIdx.x] = peer_bufs[threadIdx.x];  // immutable table
  __syncthreads();
  pdl_wait();
  // Read AFTER the wait: under programmatic dependent launch this grid may start while an earlier all-reduce of the same
  // stream is still running (every kernel of the step triggers its dependents at once), and that one updates the counters.
  const uint32_t flag = counters[blockIdx.x] + 1;
  const int64_t half_off = AR_SIGNAL_BYTES + (int64_t)(flag & 1) * WORLD * slot_bytes;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x, first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  // (A) push
  for (int64_t i = first; i < n_vec; i += stride) {
    const V v = reinterpret_cast<const V*>(in)[i];
#pragma unroll
    for (int r = 0; r < WORLD; ++r) reinterpret_cast<V*>(s_bufs[r] + half_off + rank * slot_bytes)[i] = v;
  }
  // (B) barrier: the release publishes this CTA's pushes (ordered before it by the CTA barrier) to each peer
  __syncthreads();
  if (threadIdx.x < WORLD) {
    st_release_sys(&reinterpret_cast<ArSignals*>(s_bufs[threadIdx.x])->arrived[blockIdx.x][rank], flag);
    const uint32_t* w = &reinterpret_cast<ArSignals*>(s_bufs[rank])->arrived[blockIdx.x][threadIdx.x];
    // a peer that never arrives (crashed rank, mismatched call sequence) must end as a launch failure, not as a hung GPU:
    // back off after the first microseconds and trap after about a minute
    for (uint32_t spins = 0; (int32_t)(ld_acquire_sys(w) - flag) < 0; ++spins) {
      if (spins > 4096) __nanosleep(256);
      if (spins > (1u << 28)) __trap();
    }
  }
  __syncthreads();
  // (C) local sum in rank order
  const uint8_t* mine = s_bufs[rank] + half_off;
  for (int64_t i = first; i < n_vec; i += stride) {
    V v[WORLD];
#pragma unroll
    for (int r = 0; r < WORLD; ++r) v[r] = __ldcv(reinterpret_cast<const V*>(mine + r * slot_bytes) + i);  // peers wrote them: no stale L1
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
  // every counter advances once per call, also those of CTAs this call did not need: the flag (and with it the buffer
  // half) is then the same call number in every CTA whatever grid sizes the calls used
  if (threadIdx.x == 0) counters[blockIdx.x] = flag;
  if (blockIdx.x == 0)
    for (int b = gridDim.x + threadIdx.x; b < AR_MAX_BLOCKS; b += blockDim.x) counters[b] = flag;
}

template <typename T>
int launch_one_shot(void* out, const void* in, const void* const* peer_bufs, uint32_t* counters, int rank, int world, int64_t n_vec,
                    int64_t slot_bytes, int blocks, cudaStream_t stream) {
#define APH_AR(W)                                                                                                             \
  case W:                                                                                                                     \
    APH_LAUNCH_PDL((one_shot_all_reduce_kernel<T, W>), dim3((unsigned)blocks), dim3(512), (size_t)0, stream, (T*)out, (const T*)in, \
                   (uint8_t* const*)peer_bufs, counters, rank, n_vec, slot_bytes);                                                      \
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

extern "C" size_t aph_all_reduce_buffer_bytes(int world, int64_t slot_bytes) {
  return (size_t)AR_SIGNAL_BYTES + (size_t)2 * world * slot_bytes;
}

extern "C" int aph_one_shot_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank,
                                       int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream) {
  const char* op = "one_shot_all_reduce";
  APH_CHECK(world >= 2 && world <= AR_MAX_RANKS && rank >= 0 && rank < world, APH_ERR_INVALID_ARGUMENT, "%s: rank %d of %d", op, rank,
            world);
  APH_CHECK(numel >= 0, APH_ERR_INVALID_ARGUMENT, "%s: negative size", op);
  if (numel == 0) return APH_OK;
  APH_CHECK(out && in && peer_bufs_dev && counters, APH_ERR_INVALID_ARGUMENT, "%s: null pointer", op);
  const int esz = dtype == APH_DT_F32 ? 4 : 2;
  APH_CHECK((numel * esz) % 16 == 0 && aligned16(out) && aligned16(in), APH_ERR_INVALID_ARGUMENT,
            "%s: the message must be a multiple of 16 bytes and 16-byte aligned", op);
  APH_CHECK(slot_bytes % 16 == 0 && numel * esz <= slot_bytes, APH_ERR_INVALID_ARGUMENT,
            "%s: message of %lld bytes does not fit the %lld-byte slots", op, (long long)(numel * esz), (long long)slot_bytes);
  const int64_t n_vec = numel * esz / 16;
  int blocks = (int)std::min<int64_t>((n_vec + 511) / 512, std::min(max_blocks > 0 ? max_blocks : AR_MAX_BLOCKS, AR_MAX_BLOCKS));
  blocks = std::max(blocks, 1);
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case APH_DT_F32: return launch_one_shot<float>(out, in, peer_bufs_dev, counters, rank, world, n_vec, slot_bytes, blocks, s);
    case APH_DT_F16: return launch_one_shot<__half>(out, in, peer_bufs_dev, counters, rank, world, n_vec, slot_bytes, blocks, s);
    case APH_DT_BF16: return launch_one_shot<__nv_bfloat16>(out, in, peer_bufs_dev, counters, rank, world, n_vec, slot_bytes, blocks, s);
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "%s: unknown dtype %d", op, dtype);
  }
}
