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
// TWO-SHOT variants (replace cross_device_reduce_2stage, custom_all_reduce.cuh:321-400) for larger groups, where pushing
// the whole message to every peer costs (world - 1) x the message of NVLink egress per rank (7 MiB for the 1 MiB
// messages of a TP 8 decode step): every 16-byte vector has an OWNER rank (by the thread that handles it, so the CTA
// pairing of the barrier still holds); (A) push each vector to its owner only, (B1) barrier, (C1) the owner adds the
// `world` partials in rank order in fp32, rounds once and pushes the sum into slot `owner` of every rank - that slot
// position is unused on the other ranks, (B2) barrier, (C2) every rank reads the sums from its own buffer. Egress per
// rank: 2 (world - 1) / world x the message. Same double buffering, same bit-identical results on every rank.
// The per-CTA counters hold the CALL number n; call n uses half n & 1 and the barrier flags 2n and 2n + 1, so one-shot
// and two-shot calls can be mixed freely.
// MULTIMEM variants (NVLS: the reduction happens inside the NVSwitch): the host also passes the MULTICAST mapping of the
// symmetric buffer - one address that stands for the same offset in every rank's buffer. (A) every rank copies its
// partial into slot 0 of its OWN buffer (no NVLink traffic), (B1) barrier, (C1) the owner of a vector issues ONE
// multimem.ld_reduce on the multicast address (the switch reads the `world` copies and returns their fp32-accumulated sum)
// and ONE multimem.st of the rounded sum to slot 1 (the switch writes it into every rank's buffer), (B2) barrier, (C2)
// every rank reads slot 1 of its own buffer. NVLink bytes per rank: about 1 x the message in each direction (two-shot:
// 2 (world - 1) / world), and each vector costs its owner two instructions instead of 2 x world. The order of the
// additions is the switch's, not rank order: the result is identical on all ranks (one owner computes it) and the same
// from run to run, but may differ from the rank-order sum in the last bit.
// LL variants (the low-latency protocol of NCCL's "LL" algorithm restated for this two-shot exchange): every 8 bytes on the
// wire are 4 bytes of data + a 4-byte flag = the call number, written with one store (8-byte stores are single-copy
// atomic), so DATA ARRIVAL IS THE SIGNAL: no flag barrier, no system-scope fence and no round trip anywhere. (A) a thread
// pushes each vector it does not own to the owner's receive lines, (C1) the owner polls the world - 1 lines of a vector
// until their flags show this call, adds in rank order in fp32, rounds once and pushes the sum to every rank's result
// lines, (C2) everybody polls its result lines. Twice the bytes on the wire of two-shot, but two one-way NVLink latencies
// instead of two barriers of about 4 us each: the choice up to AR_LL_MAX_BYTES. The lines live in their own region
// behind the slots (zeroed once by the host; a flag is never 0 and repeats only after 2^32 calls, and a line is only rewritten after its reader has
// moved on: a rank can push call n + 1 only after it has received every owner's results of call n, which the owners send
// after reading all lines of call n). Same rank-order sums as one-shot / two-shot: bit-identical to them.
// Layout of a rank's buffer (from the host: torch symmetric memory / CUDA IPC): [ 4 KiB flags | 2 halves x world slots
// x slot_bytes | LL lines ]. A CTA only waits for the CTA with the same index on the peers and grids are at most 128 CTAs of 512
// threads, so every waiting CTA is resident (no deadlock on a 148-SM part).
#include <type_traits>

#include "common.cuh"

namespace aphb200 {
namespace {

constexpr int AR_MAX_BLOCKS = 128;
constexpr int AR_MAX_RANKS = 8;
constexpr int AR_SIGNAL_BYTES = 4096;

// LL: messages up to 2 MiB; a "row" is the 512 vectors one pass of a CTA covers
constexpr int64_t AR_LL_MAX_BYTES = 2 << 20;
constexpr int AR_LL_ROWS = 512;
__host__ __device__ constexpr int ar_ll_chunk(int world) { return (512 + world - 1) / world; }  // threads (= vectors of a row) per owner
__host__ __device__ constexpr int64_t ar_ll_recv_bytes(int world) { return (int64_t)world * AR_LL_ROWS * ar_ll_chunk(world) * 32; }
__host__ __device__ constexpr int64_t ar_ll_result_bytes() { return (int64_t)AR_LL_ROWS * 512 * 32; }

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

// sum over the ranks of the 16 bytes at multicast address p (fp32 accumulation for the 16-bit types), and the broadcast store
template <typename T>
__device__ __forceinline__ uint4 multimem_ld_reduce(const void* p) {
  uint4 v;
  if constexpr (std::is_same<T, float>::value) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  } else if constexpr (std::is_same<T, __half>::value) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  }
  return v;
}
__device__ __forceinline__ void multimem_st(void* p, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// One 16-byte vector = two LL lines of {data, flag, data, flag}. The loads poll until both flags show this call.
__device__ __forceinline__ void ll_store(uint8_t* lines, const uint4& v, uint32_t flag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"(lines), "r"(v.x), "r"(flag), "r"(v.y) : "memory");
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %2};" ::"l"(lines + 16), "r"(v.z), "r"(flag), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ll_load(const uint8_t* lines, uint32_t flag) {
  uint4 v;
  uint32_t f0, f1, f2, f3;
  for (uint32_t spins = 0;; ++spins) {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(f0), "=r"(v.y), "=r"(f1) : "l"(lines) : "memory");
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.z), "=r"(f2), "=r"(v.w), "=r"(f3) : "l"(lines + 16) : "memory");
    if (f0 == flag && f1 == flag && f2 == flag && f3 == flag) break;
    if (spins > 4096) __nanosleep(128);
    if (spins > (1u << 27)) __trap();  // a peer that never sends: end as a launch failure, not as a hung GPU
  }
  return v;
}
// rank-order fp32 sum of one vector: `own` for this rank, the received lines for the others. All world - 1 pairs of
// lines are requested before the first flag is looked at (one L2 latency per attempt, not one per peer).
template <typename T, int WORLD>
__device__ __forceinline__ uint4 ll_sum(const uint4& own, const uint8_t* recv, int64_t src_stride, int rank, uint32_t flag) {
  constexpr int VEC = 16 / (int)sizeof(T);
  uint4 v[WORLD];
  for (uint32_t spins = 0;; ++spins) {
    uint32_t f[WORLD][4];
#pragma unroll
    for (int r = 0; r < WORLD; ++r) {
      const uint8_t* lines = recv + r * src_stride;
      if (r != rank) {
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v[r].x), "=r"(f[r][0]), "=r"(v[r].y), "=r"(f[r][1]) : "l"(lines) : "memory");
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v[r].z), "=r"(f[r][2]), "=r"(v[r].w), "=r"(f[r][3])
                     : "l"(lines + 16)
                     : "memory");
      }
    }
    bool ok = true;
#pragma unroll
    for (int r = 0; r < WORLD; ++r)
      if (r != rank) ok = ok && f[r][0] == flag && f[r][1] == flag && f[r][2] == flag && f[r][3] == flag;
    if (ok) break;
    if (spins > 4096) __nanosleep(128);
    if (spins > (1u << 27)) __trap();  // a peer that never sends: end as a launch failure, not as a hung GPU
  }
  float acc[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) acc[e] = ScalarTraits<T>::to_float(reinterpret_cast<const T*>(rank == 0 ? &own : &v[0])[e]);
#pragma unroll
  for (int r = 1; r < WORLD; ++r)
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] += ScalarTraits<T>::to_float(reinterpret_cast<const T*>(r == rank ? &own : &v[r])[e]);
  uint4 o;
#pragma unroll
  for (int e = 0; e < VEC; ++e) reinterpret_cast<T*>(&o)[e] = ScalarTraits<T>::from_float(acc[e]);
  return o;
}

// Barrier with the CTA of the same index on every peer: publish `flag` (release: this CTA's earlier stores, ordered
// before it by the CTA barrier, become visible with it) and wait until every peer has published at least `flag`.
// A peer that never arrives (crashed rank, mismatched call sequence) must end as a launch failure, not as a hung GPU:
// back off after the first microseconds and trap after about a minute.
template <int WORLD>
__device__ __forceinline__ void ar_barrier(uint8_t* const* s_bufs, int rank, uint32_t flag) {
  __syncthreads();
  if (threadIdx.x < WORLD) {
    st_release_sys(&reinterpret_cast<ArSignals*>(s_bufs[threadIdx.x])->arrived[blockIdx.x][rank], flag);
    const uint32_t* w = &reinterpret_cast<ArSignals*>(s_bufs[rank])->arrived[blockIdx.x][threadIdx.x];
    for (uint32_t spins = 0; (int32_t)(ld_acquire_sys(w) - flag) < 0; ++spins) {
      if (spins > 4096) __nanosleep(256);
      if (spins > (1u << 28)) __trap();
    }
  }
  __syncthreads();
}

// every counter advances once per call, also those of CTAs this call did not need: the call number (and with it the
// buffer half) is then the same in every CTA whatever grid sizes the calls used
__device__ __forceinline__ void ar_finish(uint32_t* counters, uint32_t call) {
  if (threadIdx.x == 0) counters[blockIdx.x] = call;
  if (blockIdx.x == 0)
    for (int b = gridDim.x + threadIdx.x; b < AR_MAX_BLOCKS; b += blockDim.x) counters[b] = call;
}

template <typename T, int WORLD>
__device__ __forceinline__ uint4 ar_sum_slots(const uint8_t* mine, int64_t slot_bytes, int64_t idx) {
  constexpr int VEC = 16 / (int)sizeof(T);
  uint4 v[WORLD];
#pragma unroll
  for (int r = 0; r < WORLD; ++r) v[r] = __ldcv(reinterpret_cast<const uint4*>(mine + r * slot_bytes) + idx);  // peers wrote them: no stale L1
  float acc[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) acc[e] = ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&v[0])[e]);
#pragma unroll
  for (int r = 1; r < WORLD; ++r)
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc[e] += ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&v[r])[e]);
  uint4 o;
#pragma unroll
  for (int e = 0; e < VEC; ++e) reinterpret_cast<T*>(&o)[e] = ScalarTraits<T>::from_float(acc[e]);
  return o;
}

constexpr int AR_ONE_SHOT = 0, AR_TWO_SHOT = 1, AR_MULTIMEM = 2, AR_LL = 3;

template <typename T, int WORLD, int MODE>
__global__ void __launch_bounds__(512, 1)
    one_shot_all_reduce_kernel(T* out, const T* in, uint8_t* const* __restrict__ peer_bufs,  // out may alias in
                               uint8_t* multicast_buf, uint32_t* __restrict__ counters, int rank, int64_t n_vec, int64_t slot_bytes) {
  using V = uint4;
  __shared__ uint8_t* s_bufs[WORLD];
  pdl_trigger();
  if (threadIdx.x < WORLD) s_bufs[threadIdx.x] = peer_bufs[threadIdx.x];  // immutable table
  __syncthreads();
  pdl_wait();
  // Read AFTER the wait: under programmatic dependent launch this grid may start while an earlier all-reduce of the same
  // stream is still running (every kernel of the step triggers its dependents at once), and that one updates the counters.
  const uint32_t call = counters[blockIdx.x] + 1;
  const uint32_t ll_flag = call ? call : 0x80000000u;  // never 0: that is what a line that was never written holds
  const int64_t half_off = AR_SIGNAL_BYTES + (int64_t)(call & 1) * WORLD * slot_bytes;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x, first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint8_t* mine = s_bufs[rank] + half_off;
  if (MODE == AR_ONE_SHOT) {
    // (A) push to everybody
    for (int64_t i = first; i < n_vec; i += stride) {
      const V v = reinterpret_cast<const V*>(in)[i];
#pragma unroll
      for (int r = 0; r < WORLD; ++r) reinterpret_cast<V*>(s_bufs[r] + half_off + rank * slot_bytes)[i] = v;
    }
    ar_barrier<WORLD>(s_bufs, rank, 2 * call);  // (B)
    // (C) local sum in rank order
    for (int64_t i = first; i < n_vec; i += stride) reinterpret_cast<V*>(out)[i] = ar_sum_slots<T, WORLD>(mine, slot_bytes, i);
  } else if (MODE == AR_TWO_SHOT) {
    const int owner = (int)threadIdx.x * WORLD / (int)blockDim.x;  // of every vector this thread handles
    // (A) push to the owner only
    for (int64_t i = first; i < n_vec; i += stride)
      reinterpret_cast<V*>(s_bufs[owner] + half_off + rank * slot_bytes)[i] = reinterpret_cast<const V*>(in)[i];
    ar_barrier<WORLD>(s_bufs, rank, 2 * call);  // (B1)
    // (C1) the owner reduces and hands the sum to everybody, in slot `rank` (unused at these positions on the peers)
    if (owner == rank)
      for (int64_t i = first; i < n_vec; i += stride) {
        const V o = ar_sum_slots<T, WORLD>(mine, slot_bytes, i);
#pragma unroll
        for (int r = 0; r < WORLD; ++r) reinterpret_cast<V*>(s_bufs[r] + half_off + rank * slot_bytes)[i] = o;
      }
    ar_barrier<WORLD>(s_bufs, rank, 2 * call + 1);  // (B2)
    // (C2)
    for (int64_t i = first; i < n_vec; i += stride) reinterpret_cast<V*>(out)[i] = __ldcv(reinterpret_cast<const V*>(mine + owner * slot_bytes) + i);
  } else if (MODE == AR_LL) {
    constexpr int CS = ar_ll_chunk(WORLD);
    const int owner = (int)threadIdx.x / CS, lj = (int)threadIdx.x - owner * CS;
    const int64_t ll_off = AR_SIGNAL_BYTES + (int64_t)2 * WORLD * slot_bytes;  // receive lines [source][row][chunk], then result lines [row][512]
    const int64_t src_stride = (int64_t)AR_LL_ROWS * CS * 32;
    // (A) vectors owned elsewhere go to their owner's receive lines
    if (owner != rank)
      for (int64_t i = first; i < n_vec; i += stride)
        ll_store(s_bufs[owner] + ll_off + rank * src_stride + ((i >> 9) * CS + lj) * 32, reinterpret_cast<const V*>(in)[i], ll_flag);
    // (C1) the owner adds in rank order as the lines arrive and sends the rounded sum to every rank's result lines
    if (owner == rank)
      for (int64_t i = first; i < n_vec; i += stride) {
        const V o = ll_sum<T, WORLD>(reinterpret_cast<const V*>(in)[i], s_bufs[rank] + ll_off + ((i >> 9) * CS + lj) * 32, src_stride, rank, ll_flag);
#pragma unroll
        for (int r = 0; r < WORLD; ++r) ll_store(s_bufs[r] + ll_off + ar_ll_recv_bytes(WORLD) + i * 32, o, ll_flag);
      }
    // (C2)
    for (int64_t i = first; i < n_vec; i += stride)
      reinterpret_cast<V*>(out)[i] = ll_load(s_bufs[rank] + ll_off + ar_ll_recv_bytes(WORLD) + i * 32, ll_flag);
  } else {
    const int owner = (int)threadIdx.x * WORLD / (int)blockDim.x;
    // (A) my partial into slot 0 of my own buffer
    for (int64_t i = first; i < n_vec; i += stride)
      reinterpret_cast<V*>(s_bufs[rank] + half_off)[i] = reinterpret_cast<const V*>(in)[i];
    ar_barrier<WORLD>(s_bufs, rank, 2 * call);  // (B1)
    // (C1) the owner lets the switch add the `world` copies and broadcast the sum into slot 1 of every buffer. The
    // multimem.st are relaxed.sys (strong) operations: the CTA barrier + release.sys flag store of ar_barrier orders them
    // before the flag like the pushes of the other variants.
    if (owner == rank) {
      uint8_t* mc = multicast_buf + half_off;
      for (int64_t i = first; i < n_vec; i += stride) multimem_st(reinterpret_cast<V*>(mc + slot_bytes) + i, multimem_ld_reduce<T>(reinterpret_cast<const V*>(mc) + i));
    }
    ar_barrier<WORLD>(s_bufs, rank, 2 * call + 1);  // (B2)
    // (C2)
    for (int64_t i = first; i < n_vec; i += stride) reinterpret_cast<V*>(out)[i] = __ldcv(reinterpret_cast<const V*>(mine + slot_bytes) + i);
  }
  ar_finish(counters, call);
}

// ---------------------------------------------------------------------------------------------------------------
// All-reduce fused with what always follows it in a decoder layer (llama.py:270-292): residual add + RMSNorm
// (fused_add_rms_norm, layernorm_kernels.cu:50-137). Same push / barrier as above, but CTAs own whole ROWS, so phase (C)
// can finish the row: s = T(sum over ranks) (the all-reduce's rounding), z = T(s + residual) -> residual, then
// x = T(z * rsqrt(mean(z^2) + eps)) * weight. One kernel and one pass over the row instead of two kernels.
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T ar_round(float v) { return ScalarTraits<T>::from_float(v); }

__device__ __forceinline__ float ar_block_sum(float v, float* s_part) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = (blockDim.x + 31) >> 5;
  __syncthreads();  // s_part may still be read from the previous row
  if (lane == 0) s_part[warp] = v;
  __syncthreads();
  v = lane < nwarp ? s_part[lane] : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T, int WORLD, int MODE>
__global__ void __launch_bounds__(512, 1)
    all_reduce_add_rms_norm_kernel(T* x, T* residual, const T* __restrict__ weight, float epsilon,
                                   uint8_t* const* __restrict__ peer_bufs, uint8_t* multicast_buf, uint32_t* __restrict__ counters,
                                   int rank, int rows, int hidden, int64_t slot_bytes) {
  constexpr bool TWO_SHOT = MODE == AR_TWO_SHOT, MULTIMEM = MODE == AR_MULTIMEM, LL = MODE == AR_LL;
  constexpr int CS = ar_ll_chunk(WORLD);
  constexpr int VEC = 16 / (int)sizeof(T);
  constexpr int ITER = 4;  // hidden <= 4 * 512 * VEC
  using V = uint4;
  __shared__ uint8_t* s_bufs[WORLD];
  __shared__ float s_part[32];
  pdl_trigger();
  if (threadIdx.x < WORLD) s_bufs[threadIdx.x] = peer_bufs[threadIdx.x];
  const int nvec = hidden / VEC;
  V wreg[ITER];  // immutable: before the wait
#pragma unroll
  for (int it = 0; it < ITER; ++it) {
    const int i = threadIdx.x + it * blockDim.x;
    if (i < nvec) wreg[it] = reinterpret_cast<const V*>(weight)[i];
  }
  __syncthreads();
  pdl_wait();
  const uint32_t call = counters[blockIdx.x] + 1;  // after the wait, see one_shot_all_reduce_kernel
  const uint32_t ll_flag = call ? call : 0x80000000u;  // never 0: that is what a line that was never written holds
  const int64_t half_off = AR_SIGNAL_BYTES + (int64_t)(call & 1) * WORLD * slot_bytes;
  const uint8_t* mine = s_bufs[rank] + half_off;
  // two-shot, multimem, LL: the owner of every vector this thread handles
  const int owner = LL ? (int)threadIdx.x / CS : (int)threadIdx.x * WORLD / (int)blockDim.x;
  const int lj = (int)threadIdx.x - owner * CS;
  const int64_t ll_off = AR_SIGNAL_BYTES + (int64_t)2 * WORLD * slot_bytes, src_stride = (int64_t)AR_LL_ROWS * CS * 32;
  const int passes = (nvec + (int)blockDim.x - 1) / (int)blockDim.x;  // LL: line row of (row, pass) = row * passes + pass < AR_LL_ROWS
  if (LL) {
    // (A) + (C1) without barriers: push what others own, add what I own as it arrives and hand the sums to every rank
    for (int row = blockIdx.x; row < rows; row += gridDim.x)
      for (int i = threadIdx.x, ps = 0; i < nvec; i += blockDim.x, ++ps)
        if (owner != rank)
          ll_store(s_bufs[owner] + ll_off + rank * src_stride + ((int64_t)(row * passes + ps) * CS + lj) * 32,
                   reinterpret_cast<const V*>(x)[(int64_t)row * nvec + i], ll_flag);
    if (owner == rank)
      for (int row = blockIdx.x; row < rows; row += gridDim.x)
        for (int i = threadIdx.x, ps = 0; i < nvec; i += blockDim.x, ++ps) {
          const int64_t lrow = (int64_t)row * passes + ps;
          const V o = ll_sum<T, WORLD>(reinterpret_cast<const V*>(x)[(int64_t)row * nvec + i], s_bufs[rank] + ll_off + (lrow * CS + lj) * 32,
                                       src_stride, rank, ll_flag);
#pragma unroll
          for (int r = 0; r < WORLD; ++r) ll_store(s_bufs[r] + ll_off + ar_ll_recv_bytes(WORLD) + (lrow * 512 + threadIdx.x) * 32, o, ll_flag);
        }
  }
  // (A) push this CTA's rows: to everybody, or (two-shot) to the owner of each vector
  if (!LL)
  for (int row = blockIdx.x; row < rows; row += gridDim.x)
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      const int64_t idx = (int64_t)row * nvec + i;
      const V v = reinterpret_cast<const V*>(x)[idx];
      if (MULTIMEM) {
        reinterpret_cast<V*>(s_bufs[rank] + half_off)[idx] = v;  // slot 0 of my own buffer
      } else if (TWO_SHOT) {
        reinterpret_cast<V*>(s_bufs[owner] + half_off + rank * slot_bytes)[idx] = v;
      } else {
#pragma unroll
        for (int r = 0; r < WORLD; ++r) reinterpret_cast<V*>(s_bufs[r] + half_off + rank * slot_bytes)[idx] = v;
      }
    }
  if (!LL) ar_barrier<WORLD>(s_bufs, rank, 2 * call);  // (B) / (B1)
  if (TWO_SHOT) {
    // (C1) the owner reduces its vectors and hands the rounded sums to every rank (slot `rank` at these positions)
    if (owner == rank)
      for (int row = blockIdx.x; row < rows; row += gridDim.x)
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
          const int64_t idx = (int64_t)row * nvec + i;
          const V o = ar_sum_slots<T, WORLD>(mine, slot_bytes, idx);
#pragma unroll
          for (int r = 0; r < WORLD; ++r) reinterpret_cast<V*>(s_bufs[r] + half_off + rank * slot_bytes)[idx] = o;
        }
    ar_barrier<WORLD>(s_bufs, rank, 2 * call + 1);  // (B2)
  }
  if (MULTIMEM) {
    // (C1) the owner lets the switch add the copies and broadcast the rounded sums into slot 1 of every buffer
    if (owner == rank) {
      uint8_t* mc = multicast_buf + half_off;
      for (int row = blockIdx.x; row < rows; row += gridDim.x)
        for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
          const int64_t idx = (int64_t)row * nvec + i;
          multimem_st(reinterpret_cast<V*>(mc + slot_bytes) + idx, multimem_ld_reduce<T>(reinterpret_cast<const V*>(mc) + idx));
        }
    }
    ar_barrier<WORLD>(s_bufs, rank, 2 * call + 1);  // (B2)
  }
  // (C) finish the rows
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    V keep[ITER];
    float ss = 0.f;
#pragma unroll
    for (int it = 0; it < ITER; ++it) {
      const int i = threadIdx.x + it * blockDim.x;
      if (i < nvec) {
        const int64_t idx = (int64_t)row * nvec + i;
        // the all-reduce's output, rounded to T: summed here in rank order (one-shot) or by the vector's owner (two-shot)
        const V sum = LL         ? ll_load(s_bufs[rank] + ll_off + ar_ll_recv_bytes(WORLD) + (((int64_t)row * passes + it) * 512 + threadIdx.x) * 32, ll_flag)
                      : MULTIMEM ? __ldcv(reinterpret_cast<const V*>(mine + slot_bytes) + idx)
                      : TWO_SHOT ? __ldcv(reinterpret_cast<const V*>(mine + owner * slot_bytes) + idx)
                                 : ar_sum_slots<T, WORLD>(mine, slot_bytes, idx);
        const V res = reinterpret_cast<const V*>(residual)[idx];
        V z;
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
          const float s = ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&sum)[e]);                     // all-reduce output
          const T zz = ar_round<T>(s + ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&res)[e]));       // scalar_t add
          reinterpret_cast<T*>(&z)[e] = zz;
          const float f = ScalarTraits<T>::to_float(zz);
          ss += f * f;
        }
        reinterpret_cast<V*>(residual)[idx] = z;
        keep[it] = z;
      }
    }
    ss = ar_block_sum(ss, s_part);
    const float rs = rsqrtf(ss / (float)hidden + epsilon);
#pragma unroll
    for (int it = 0; it < ITER; ++it) {
      const int i = threadIdx.x + it * blockDim.x;
      if (i < nvec) {
        V o;
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
          const T n = ar_round<T>(ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&keep[it])[e]) * rs);
          reinterpret_cast<T*>(&o)[e] =
              ar_round<T>(ScalarTraits<T>::to_float(n) * ScalarTraits<T>::to_float(reinterpret_cast<const T*>(&wreg[it])[e]));
        }
        reinterpret_cast<V*>(x)[(int64_t)row * nvec + i] = o;
      }
    }
  }
  ar_finish(counters, call);
}

template <typename T>
int launch_ar_norm(void* x, void* residual, const void* weight, float eps, const void* const* peer_bufs, void* mc, uint32_t* counters,
                   int rank, int world, int rows, int hidden, int64_t slot_bytes, int mode, cudaStream_t stream) {
  const int blocks = std::max(1, std::min(rows, AR_MAX_BLOCKS));
#define APH_ARN_MODE(W, MODE)                                                                                                      \
  APH_LAUNCH_PDL((all_reduce_add_rms_norm_kernel<T, W, MODE>), dim3((unsigned)blocks), dim3(512), (size_t)0, stream, (T*)x,        \
                 (T*)residual, (const T*)weight, eps, (uint8_t* const*)peer_bufs, (uint8_t*)mc, counters, rank, rows, hidden,       \
                 slot_bytes)
#define APH_ARN(W)                        \
  case W:                                 \
    if (mode == AR_TWO_SHOT) {            \
      APH_ARN_MODE(W, AR_TWO_SHOT);       \
    } else if (mode == AR_MULTIMEM) {     \
      APH_ARN_MODE(W, AR_MULTIMEM);       \
    } else if (mode == AR_LL) {           \
      APH_ARN_MODE(W, AR_LL);             \
    } else {                              \
      APH_ARN_MODE(W, AR_ONE_SHOT);       \
    }                                     \
    return APH_OK
  switch (world) {
    APH_ARN(2);
    APH_ARN(4);
    APH_ARN(6);
    APH_ARN(8);
    default: return set_error(APH_ERR_UNSUPPORTED, "all_reduce_add_rms_norm: world size %d (2, 4, 6 or 8)", world);
  }
#undef APH_ARN
#undef APH_ARN_MODE
}

template <typename T>
int launch_one_shot(void* out, const void* in, const void* const* peer_bufs, void* mc, uint32_t* counters, int rank, int world,
                    int64_t n_vec, int64_t slot_bytes, int blocks, int mode, cudaStream_t stream) {
#define APH_AR_MODE(W, MODE)                                                                                                    \
  APH_LAUNCH_PDL((one_shot_all_reduce_kernel<T, W, MODE>), dim3((unsigned)blocks), dim3(512), (size_t)0, stream, (T*)out,      \
                 (const T*)in, (uint8_t* const*)peer_bufs, (uint8_t*)mc, counters, rank, n_vec, slot_bytes)
#define APH_AR(W)                        \
  case W:                                \
    if (mode == AR_TWO_SHOT) {           \
      APH_AR_MODE(W, AR_TWO_SHOT);       \
    } else if (mode == AR_MULTIMEM) {    \
      APH_AR_MODE(W, AR_MULTIMEM);       \
    } else if (mode == AR_LL) {          \
      APH_AR_MODE(W, AR_LL);             \
    } else {                             \
      APH_AR_MODE(W, AR_ONE_SHOT);       \
    }                                    \
    return APH_OK
  switch (world) {
    APH_AR(2);
    APH_AR(4);
    APH_AR(6);
    APH_AR(8);
    default: return set_error(APH_ERR_UNSUPPORTED, "all_reduce: world size %d (2, 4, 6 or 8)", world);
  }
#undef APH_AR
#undef APH_AR_MODE
}

int all_reduce_entry(const char* op, int mode, void* out, const void* in, const void* const* peer_bufs_dev, void* mc, uint32_t* counters,
                     int rank, int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream) {
  APH_CHECK(mode != AR_MULTIMEM || (mc != nullptr && aligned16(mc)), APH_ERR_INVALID_ARGUMENT, "%s: null or misaligned multicast address", op);
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
  APH_CHECK(mode != AR_LL || numel * esz <= AR_LL_MAX_BYTES, APH_ERR_INVALID_ARGUMENT, "%s: the LL lines hold messages up to %lld bytes", op,
            (long long)AR_LL_MAX_BYTES);
  const int64_t n_vec = numel * esz / 16;
  int blocks = (int)std::min<int64_t>((n_vec + 511) / 512, std::min(max_blocks > 0 ? max_blocks : AR_MAX_BLOCKS, AR_MAX_BLOCKS));
  blocks = std::max(blocks, 1);
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case APH_DT_F32: return launch_one_shot<float>(out, in, peer_bufs_dev, mc, counters, rank, world, n_vec, slot_bytes, blocks, mode, s);
    case APH_DT_F16: return launch_one_shot<__half>(out, in, peer_bufs_dev, mc, counters, rank, world, n_vec, slot_bytes, blocks, mode, s);
    case APH_DT_BF16:
      return launch_one_shot<__nv_bfloat16>(out, in, peer_bufs_dev, mc, counters, rank, world, n_vec, slot_bytes, blocks, mode, s);
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "%s: unknown dtype %d", op, dtype);
  }
}

int ar_norm_entry(const char* op, int mode, void* x, void* residual, const void* weight, float epsilon, const void* const* peer_bufs_dev,
                  void* mc, uint32_t* counters, int rank, int world, int64_t num_tokens, int64_t hidden_size, int dtype, int64_t slot_bytes,
                  void* stream) {
  APH_CHECK(mode != AR_MULTIMEM || (mc != nullptr && aligned16(mc)), APH_ERR_INVALID_ARGUMENT, "%s: null or misaligned multicast address", op);
  APH_CHECK(world >= 2 && world <= AR_MAX_RANKS && rank >= 0 && rank < world, APH_ERR_INVALID_ARGUMENT, "%s: rank %d of %d", op, rank,
            world);
  APH_CHECK(dtype == APH_DT_F16 || dtype == APH_DT_BF16, APH_ERR_UNSUPPORTED, "%s: float16 / bfloat16 only", op);
  APH_CHECK(num_tokens >= 0 && hidden_size > 0 && hidden_size % 8 == 0 && hidden_size <= 4 * 512 * 8, APH_ERR_INVALID_ARGUMENT,
            "%s: hidden size must be a multiple of 8 and at most 16384", op);
  if (num_tokens == 0) return APH_OK;
  APH_CHECK(x && residual && weight && peer_bufs_dev && counters, APH_ERR_INVALID_ARGUMENT, "%s: null pointer", op);
  APH_CHECK(aligned16(x) && aligned16(residual) && aligned16(weight), APH_ERR_INVALID_ARGUMENT, "%s: tensors must be 16-byte aligned", op);
  APH_CHECK(slot_bytes % 16 == 0 && num_tokens * hidden_size * 2 <= slot_bytes, APH_ERR_INVALID_ARGUMENT,
            "%s: message does not fit the %lld-byte slots", op, (long long)slot_bytes);
  APH_CHECK(mode != AR_LL || (num_tokens * hidden_size * 2 <= AR_LL_MAX_BYTES && num_tokens * ((hidden_size / 8 + 511) / 512) <= AR_LL_ROWS),
            APH_ERR_INVALID_ARGUMENT, "%s: the LL lines hold messages up to %lld bytes in at most %d passes of 512 vectors", op,
            (long long)AR_LL_MAX_BYTES, AR_LL_ROWS);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == APH_DT_F16)
    return launch_ar_norm<__half>(x, residual, weight, epsilon, peer_bufs_dev, mc, counters, rank, world, (int)num_tokens, (int)hidden_size,
                                  slot_bytes, mode, s);
  return launch_ar_norm<__nv_bfloat16>(x, residual, weight, epsilon, peer_bufs_dev, mc, counters, rank, world, (int)num_tokens,
                                       (int)hidden_size, slot_bytes, mode, s);
}

}  // namespace
}  // namespace aphb200

using namespace aphb200;

extern "C" size_t aph_all_reduce_signal_bytes(void) { return AR_SIGNAL_BYTES; }
extern "C" int aph_all_reduce_max_blocks(void) { return AR_MAX_BLOCKS; }

extern "C" size_t aph_all_reduce_buffer_bytes(int world, int64_t slot_bytes) {
  return (size_t)AR_SIGNAL_BYTES + (size_t)2 * world * slot_bytes + (size_t)ar_ll_recv_bytes(world) + (size_t)ar_ll_result_bytes();
}
extern "C" size_t aph_all_reduce_ll_max_bytes(void) { return (size_t)AR_LL_MAX_BYTES; }
extern "C" int aph_all_reduce_ll_max_passes(void) { return AR_LL_ROWS; }

extern "C" int aph_ll_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank, int world,
                                 int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream) {
  return all_reduce_entry("ll_all_reduce", AR_LL, out, in, peer_bufs_dev, nullptr, counters, rank, world, numel, dtype, slot_bytes, max_blocks,
                          stream);
}

extern "C" int aph_ll_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon, const void* const* peer_bufs_dev,
                                              uint32_t* counters, int rank, int world, int64_t num_tokens, int64_t hidden_size, int dtype,
                                              int64_t slot_bytes, void* stream) {
  return ar_norm_entry("ll_all_reduce_add_rms_norm", AR_LL, x, residual, weight, epsilon, peer_bufs_dev, nullptr, counters, rank, world,
                       num_tokens, hidden_size, dtype, slot_bytes, stream);
}

extern "C" int aph_one_shot_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank,
                                       int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream) {
  return all_reduce_entry("one_shot_all_reduce", AR_ONE_SHOT, out, in, peer_bufs_dev, nullptr, counters, rank, world, numel, dtype, slot_bytes,
                          max_blocks, stream);
}

extern "C" int aph_two_shot_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, uint32_t* counters, int rank,
                                       int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream) {
  return all_reduce_entry("two_shot_all_reduce", AR_TWO_SHOT, out, in, peer_bufs_dev, nullptr, counters, rank, world, numel, dtype, slot_bytes,
                          max_blocks, stream);
}

extern "C" int aph_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon,
                                           const void* const* peer_bufs_dev, uint32_t* counters, int rank, int world,
                                           int64_t num_tokens, int64_t hidden_size, int dtype, int64_t slot_bytes, void* stream) {
  return ar_norm_entry("all_reduce_add_rms_norm", AR_ONE_SHOT, x, residual, weight, epsilon, peer_bufs_dev, nullptr, counters, rank, world,
                       num_tokens, hidden_size, dtype, slot_bytes, stream);
}

extern "C" int aph_two_shot_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon,
                                                    const void* const* peer_bufs_dev, uint32_t* counters, int rank, int world,
                                                    int64_t num_tokens, int64_t hidden_size, int dtype, int64_t slot_bytes,
                                                    void* stream) {
  return ar_norm_entry("two_shot_all_reduce_add_rms_norm", AR_TWO_SHOT, x, residual, weight, epsilon, peer_bufs_dev, nullptr, counters, rank,
                       world, num_tokens, hidden_size, dtype, slot_bytes, stream);
}

extern "C" int aph_multimem_all_reduce(void* out, const void* in, const void* const* peer_bufs_dev, void* multicast_buf, uint32_t* counters,
                                       int rank, int world, int64_t numel, int dtype, int64_t slot_bytes, int max_blocks, void* stream) {
  return all_reduce_entry("multimem_all_reduce", AR_MULTIMEM, out, in, peer_bufs_dev, multicast_buf, counters, rank, world, numel, dtype,
                          slot_bytes, max_blocks, stream);
}

extern "C" int aph_multimem_all_reduce_add_rms_norm(void* x, void* residual, const void* weight, float epsilon,
                                                    const void* const* peer_bufs_dev, void* multicast_buf, uint32_t* counters, int rank,
                                                    int world, int64_t num_tokens, int64_t hidden_size, int dtype, int64_t slot_bytes,
                                                    void* stream) {
  return ar_norm_entry("multimem_all_reduce_add_rms_norm", AR_MULTIMEM, x, residual, weight, epsilon, peer_bufs_dev, multicast_buf, counters,
                       rank, world, num_tokens, hidden_size, dtype, slot_bytes, stream);
}
