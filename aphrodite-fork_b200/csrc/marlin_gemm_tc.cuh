// W4A16 dequant-GEMM on Marlin-format weights, tcgen05 / tensor-memory path (sm_100a).
//
// One CTA computes a 128-column slice of the output for up to MB batch rows over a chunk of K:
//   D[n, m] (TMEM, fp32, 128 lanes x MB columns) += W[n, k] (TMEM, 16-bit) * X[m, k] (shared memory)
// i.e. the WEIGHT tile is the M=128 operand of tcgen05.mma and the batch is its N dimension, so one
// instruction shape serves every batch size from 16 to 256 and the dequantisation work per weight is
// independent of the batch.
//
//   warp 0      weight producer: one thread issues tiled TMA loads of packed weights, group scales and
//               zero points into a deep ring (8 KiB of weights per stage, released by the dequantisers
//               as soon as the words are in registers, so many stages stay in flight);
//   warp 10     activation producer: one thread loads the [MB x 128 k] activation tile with the 128-byte
//               swizzle the MMA wants into a short ring that is paired with the TMEM weight slots;
//   warps 2-9   dequantisers: each reads packed int32 words of the Marlin layout from shared memory,
//               turns one word into the four 16-bit pair registers of an m16k16 fragment (exponent
//               splice, subtract zero point, multiply by the group scale - the same roundings as the
//               reference's dequant + __hmul2) and stores four k16 steps with ONE tcgen05.st.16x128b.x8,
//               whose lane/column pattern is exactly the fragment order the Marlin format was built for;
//   warp 1      MMA issuer: one thread issues 8 tcgen05.mma (K = 16 each) per stage with A from TMEM
//               and B through a shared-memory descriptor, then tcgen05.commit releases the slot;
//   epilogue    the dequantiser warps read D with tcgen05.ld (thread = output column), apply the
//               channelwise scale if any, and write C or the split-K partial.
#pragma once
#include <type_traits>

#include "common.cuh"
#include "ptx.cuh"

namespace aphb200 {
namespace mgtc {

using namespace ::aphb200::ptx;

constexpr int NTILE = 128;    // output columns per tile (= MMA M)
constexpr int KSTAGE = 128;   // k per pipeline stage (= 8 k16-rows of the Marlin layout)
constexpr int DQ_GROUP = 8;   // dequantiser warps working on one stage: two per TMEM lane quadrant
constexpr int DQ_GROUPS = 2;  // group (it & 1) takes stage it
constexpr int W_PROD = 0, W_MMA = 1, W_ACT = 2, W_MMA2 = 3, W_EPI0 = 4, W_DQ0 = 8;  // warp roles
constexpr int EPI_WARPS = 4;
constexpr int THREADS = (W_DQ0 + DQ_GROUP * DQ_GROUPS) * 32;
constexpr int S_BYTES = 4 * NTILE * 2;
// one weight stage = 8 k16-rows x 128 columns of packed weights (4 bit: 8 KiB, 8 bit: 16 KiB) + up to 4 groups of
// scales + their zero points (packed 32 / BITS per int32)
template <int BITS>
struct WStage {
  static constexpr int PACK = 32 / BITS;
  static constexpr int ROW_BYTES = NTILE * 2 * BITS;       // one k16-row of the slice: 16 x 128 weights
  static constexpr int W_BYTES = 8 * ROW_BYTES;
  static constexpr int Z_BYTES = BITS == 4 ? S_BYTES : 4 * (NTILE / PACK) * 4;  // 4 bit: room for HQQ's 16-bit float zero points
  static constexpr int BYTES = W_BYTES + S_BYTES + Z_BYTES;
};

// The work is the flattened list of (tile, k stage) pairs, tile = (column slice, batch tile); CTA b of G
// takes the contiguous range [b*T/G, (b+1)*T/G) (stream-K): every SM gets the same number of stages
// whatever the shape. A tile whose k range is shared by several CTAs is finished by the last of them.
struct Params {
  const uint8_t* s;   // scales (channelwise path reads them directly)
  const uint8_t* zp;  // zero points or null
  void* c;
  float* scratch;     // [G][2][MB*128] fp32 partial tiles (first / last segment of each CTA)
  int* counters;      // [tiles] arrival counters, all-zero on entry and on exit
  int N, K, M;
  int group_size;     // == K for channelwise
  int k_stages;       // ceil(K / 128)
  int m_tiles;        // ceil(M / MB)
  int total;          // T = tiles * k_stages
  int cluster_split;  // 0: stream-K ranges. S >= 2: CTA b = tile * S + r is rank r of a thread-block cluster of S CTAs that
                      // share tile b / S; r takes the r-th S-th of the tile's k stages and the partials meet in shared memory
  unsigned long long* trace;  // development: per-stage clock64 stamps of CTA 0 (APH_GEMM_TRACE), else null
};

template <typename scalar_t, int MB, int BITS = 4>
struct Cfg {
  static constexpr bool F16 = std::is_same<scalar_t, __half>::value;
  static constexpr int W_BYTES = WStage<BITS>::W_BYTES, Z_BYTES = WStage<BITS>::Z_BYTES, W_STAGE = WStage<BITS>::BYTES;
  static constexpr int ACT_BOX = MB * 128;  // MB rows x 64 k, 128B-swizzled
  static constexpr int ACT_BYTES = 2 * ACT_BOX;
  static constexpr int TMEM_COLS = 512;
  static constexpr int NDBUF = MB <= 128 ? 2 : 1;  // accumulator buffers in TMEM (epilogue of one tile overlaps the next)
  // One thread sustains about one tcgen05.mma per 50-100 cycles whatever its shape (tools/micro/umma_rate.cu), which
  // is below the tensor pipe's rate for N <= 64: two issuer warps then take alternate stages, each accumulating into
  // its own D tile; the epilogue adds the two.
  static constexpr int NI = MB <= 64 ? 2 : 1;
  static constexpr int A_COL0 = NI * NDBUF * MB < 64 ? 64 : NI * NDBUF * MB;
  // activation ring == TMEM weight-slot ring
  static constexpr int SA_TMEM = (TMEM_COLS - A_COL0) / 64;
  // the partial sums of a shared tile are merged at the end of the kernel through the then idle rings
  static constexpr int FIX_BYTES = 0;
  static constexpr int BUDGET = 227 * 1024 - 1024 - 1024 - FIX_BYTES - 4096;
  static constexpr int SA_SMEM = (BUDGET - 3 * W_STAGE) / ACT_BYTES;
  // Both rings are shared by the two dequantiser groups (stage `it` goes to group it & 1) and are walked with
  // ptx::RingCursor: each group owns its own sub-ring of slots, so any depth >= 2 is safe, odd ones included (the
  // natural fit at 256 batch rows is 3 + 3). Round 1 mapped slot = it % depth and had to keep the depth even; the
  // reason (mbarrier parity aliasing when a slot alternates between the groups) is written up at RingCursor.
  static constexpr int SA_FIT = SA_SMEM < SA_TMEM ? SA_SMEM : SA_TMEM;
  static constexpr int SA = SA_FIT;
  static_assert(SA >= 2 && SA <= 7, "activation ring depth");
  static constexpr int SW_FIT = (BUDGET - SA * ACT_BYTES) / W_STAGE;
  static constexpr int SW = SW_FIT > 8 ? 8 : SW_FIT;
  static_assert(SW >= 3, "weight ring too shallow");
  static_assert(DQ_GROUPS == 2 && (NI == 1 || NI == 2), "RingCursor splits a ring between two groups taking alternate stages");
  static constexpr int ACT_OFF = 0;  // activation slots first: 1 KiB aligned
  static constexpr int W_OFF = SA * ACT_BYTES;
  static constexpr int FIX_OFF = W_OFF + SW * W_STAGE;
  static constexpr int BAR_OFF = FIX_OFF + FIX_BYTES;
  static constexpr int BAR_BYTES = 1024;
  static constexpr int EPI_STAGE = 16 * NTILE * 2;  // [16 batch rows][128 columns] 16-bit: C rows leave as 16-byte vectors
  static constexpr int STAGE_OFF = BAR_OFF + BAR_BYTES;
  static constexpr int SMEM_BYTES = 1024 + STAGE_OFF + EPI_STAGE;
  static constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;  // 1024 / 128 as packed pairs
  static constexpr uint32_t IDESC = umma_idesc_f16(F16 ? 0 : 1, NTILE, MB);
};

// first flattened index of CTA b, and the CTA that owns flattened index x
__device__ __forceinline__ int range_lo(int b, int T, int G) { return (int)(((int64_t)b * T) / G); }
__device__ __forceinline__ int owner_of(int x, int T, int G) { return (int)((((int64_t)x + 1) * G + T - 1) / T) - 1; }

template <bool F16>
__device__ __forceinline__ uint32_t sub_mul2(uint32_t x, uint32_t c, uint32_t s, bool mul) {
  if constexpr (F16) {
    __half2 v = __hsub2(*reinterpret_cast<__half2*>(&x), *reinterpret_cast<__half2*>(&c));
    if (mul) v = __hmul2(v, *reinterpret_cast<__half2*>(&s));
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __nv_bfloat162 v = __hsub2(*reinterpret_cast<__nv_bfloat162*>(&x), *reinterpret_cast<__nv_bfloat162*>(&c));
    if (mul) v = __hmul2(v, *reinterpret_cast<__nv_bfloat162*>(&s));
    return *reinterpret_cast<uint32_t*>(&v);
  }
}


// Segments of a CTA's range [lo, hi): one per tile touched. Visit order: the tail segment first when it is a
// partial tile, then the rest in natural order - so that tiles shared with a neighbour CTA are finished (and
// their partial sums merged) early, and the kernel ends on tiles this CTA owns alone.
struct SegIter {
  int lo, hi, KS, m_tiles, nseg, tile0;
  bool tail_first;
  int seg;                              // visit index
  int j, tile, k0, k1, slice, mt;       // current segment (j = natural index)
  __device__ __forceinline__ SegIter(int lo_, int hi_, int ks, int mtiles) : lo(lo_), hi(hi_), KS(ks), m_tiles(mtiles), seg(0) {
    tile0 = lo / KS;
    nseg = (hi - 1) / KS - tile0 + 1;
    tail_first = false;  // (tail-first visiting measured slower: the merge then competes with the main loop)
    set();
  }
  __device__ __forceinline__ void set() {
    j = tail_first ? (seg == 0 ? nseg - 1 : seg - 1) : seg;
    tile = tile0 + j;
    k0 = (j == 0) ? lo - tile0 * KS : 0;
    k1 = (j == nseg - 1) ? hi - tile * KS : KS;
    slice = tile / m_tiles;
    mt = tile - slice * m_tiles;
  }
  __device__ __forceinline__ bool valid() const { return seg < nseg; }
  __device__ __forceinline__ void advance() {
    ++seg;
    if (seg < nseg) set();
  }
};

// ZPF: the zero points are 16-bit FLOATS laid out like the scales (HQQ, gptq_marlin.cu:1190-1300): w = (q - z) * s
// G128: the build for group_size == 128 == one group per stage (the C3 / C4 case): the grouping tests of the generic
// build (grouped? how many groups per stage? which rows start one?) are compile-time facts, and the dequantisers'
// arithmetic becomes one straight block the scheduler can interleave freely
template <typename scalar_t, int MB, bool HAS_ZP, int BITS = 4, bool ZPF = false, bool G128 = false>
__global__ void __launch_bounds__(THREADS, 1)
    marlin_w4a16_tc_kernel(const Params p, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmA,
                           const __grid_constant__ CUtensorMap tmS, const __grid_constant__ CUtensorMap tmZ) {
  using C = Cfg<scalar_t, MB, BITS>;
  constexpr bool F16 = C::F16;
  constexpr int SA = C::SA, SW = C::SW, NDBUF = C::NDBUF;
  constexpr int W_BYTES = C::W_BYTES, W_STAGE = C::W_STAGE, PACK = 32 / BITS, ROW_BYTES = WStage<BITS>::ROW_BYTES;
  extern __shared__ uint8_t smem_raw[];

  const int tid = threadIdx.x, lane = tid & 31;
#ifdef APH_GEMM_TRACE
  if (p.trace != nullptr && blockIdx.x == 0 && tid == 0) p.trace[512 + 48 + 3] = clock64();
#endif
  // the warp index through a shuffle: ptxas then knows the role branches below are warp-uniform
  const int w = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const bool grouped = G128 || p.group_size < p.K;
  const int gps = G128 ? 1 : grouped ? KSTAGE / p.group_size : 0;  // groups per stage (1, 2, 4)
  const int G = gridDim.x, T = p.total, KS = p.k_stages;
  const int CS = p.cluster_split;
  // stream-K: equal contiguous ranges of the flattened (tile, stage) list. Cluster split-K: the r-th part of one tile.
  const int lo = CS ? (int)(blockIdx.x / CS) * KS + (int)((blockIdx.x % CS) * KS / CS) : range_lo(blockIdx.x, T, G);
  const int hi = CS ? (int)(blockIdx.x / CS) * KS + (int)((blockIdx.x % CS + 1) * KS / CS) : range_lo(blockIdx.x + 1, T, G);  // hi > lo
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar0 = smem_base + C::BAR_OFF;
  // ready[slot]: 8 dequantiser warps + the activation producer's expect_tx arrival + the bytes of its two TMA loads
  const uint32_t fullw_bar = bar0, emptyw_bar = bar0 + 64, fix_bar = bar0 + 128, ready_bar = bar0 + 192, emptya_bar = bar0 + 256,
                 done_bar = bar0 + 320, dfree_bar = bar0 + 336;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + C::BAR_OFF + 352);
  int* s_last = reinterpret_cast<int*>(smem + C::BAR_OFF + 356);

  // barrier init, one barrier per thread of the last warp (warp 1 is busy allocating TMEM meanwhile)
  if (tid >= THREADS - 32) {
    const int i = tid - (THREADS - 32);
    if (i < SW) mbar_init(fullw_bar + i * 8, 1);
    if (i >= 8 && i < 8 + SW) mbar_init(emptyw_bar + (i - 8) * 8, DQ_GROUP);
    if (i >= 16 && i < 16 + SA) mbar_init(ready_bar + (i - 16) * 8, DQ_GROUP + 1);
    if (i >= 24 && i < 24 + SA) mbar_init(emptya_bar + (i - 24) * 8, 1);
    if (i == 31) {
      mbar_init(fix_bar, 1);
      for (int s = 0; s < NDBUF; ++s) {
        mbar_init(done_bar + s * 8, C::NI);
        mbar_init(dfree_bar + s * 8, EPI_WARPS);
      }
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (tid == 0) {
    prefetch_tensormap(&tmB);
    prefetch_tensormap(&tmA);
    prefetch_tensormap(&tmS);
    prefetch_tensormap(&tmZ);
  }
  if (w == W_MMA) tmem_alloc(smem_u32(tmem_slot), C::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // programmatic dependent launch: the next kernel of the stream may be set up from here on; this kernel streams
  // its (immutable) weights at once, every role that touches activations or global outputs waits for the previous
  // kernel first (griddep_wait)
  griddep_launch_dependents();

  // Every role walks the same list of segments (tile, k0..k1) and keeps the same global stage counter `it`.
#define APH_SEGMENTS                                                                                         \
  for (SegIter sg(lo, hi, KS, p.m_tiles); sg.valid(); sg.advance())
#define APH_SEG_VARS                                                                                          \
  const int tile = sg.tile, k0 = sg.k0, k1 = sg.k1, slice = sg.slice, mt = sg.mt, seg = sg.seg, segj = sg.j;  \
  (void)tile; (void)slice; (void)mt; (void)seg; (void)segj;

  // ---- merging a shared tile: sum the partial tiles of every contributor (in CTA order: deterministic) and write C.
  //      Runs at the end of the kernel on the epilogue + dequantiser warps; the partials are pulled into the idle rings
  //      with bulk copies (full L2 bandwidth, no register round trips). ----
  constexpr int MT = (EPI_WARPS + DQ_GROUP * DQ_GROUPS) * 32;
  int* s_pend = reinterpret_cast<int*>(smem + C::BAR_OFF + 368);  // npend, tile, tile
  uint32_t fix_phase = 0;
  auto merge_tile = [&](int tile, int mtid) {
    const int slice = tile / p.m_tiles, mt = tile - slice * p.m_tiles;
    const int m_rows = min(MB, p.M - mt * MB);
    const int b_first = owner_of(tile * KS, T, G), b_last = owner_of(tile * KS + KS - 1, T, G);
    const int n_c = b_last - b_first + 1;
    scalar_t* crow = reinterpret_cast<scalar_t*>(p.c) + (int64_t)mt * MB * p.N + (int64_t)slice * NTILE;
    const int units = m_rows * (NTILE / 4);              // float4 units of the tile
    const int per = min(units, (C::BAR_OFF / 16) / n_c);  // units per contributor and chunk
    for (int u0 = 0; u0 < units; u0 += per) {
      const int cnt = min(per, units - u0);
      if (mtid == 0) {
        __threadfence();
        asm volatile("fence.proxy.async;" ::: "memory");
        mbar_expect_tx(fix_bar, (uint32_t)(n_c * cnt * 16));
        for (int j = 0; j < n_c; ++j) {
          const int bb = b_first + j;
          const int slot = (range_lo(bb, T, G) / KS == tile) ? 0 : 1;
          const float* src = p.scratch + ((size_t)bb * 2 + slot) * (size_t)(MB * NTILE) + (size_t)u0 * 4;
          bulk_g2s(smem_base + (uint32_t)(j * per * 16), src, (uint32_t)(cnt * 16), fix_bar);
        }
      }
      mbar_wait(fix_bar, fix_phase);
      fix_phase ^= 1;
      for (int idx = mtid; idx < cnt; idx += MT) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < n_c; ++j) {
          const float4 t = *reinterpret_cast<const float4*>(smem + (size_t)(j * per + idx) * 16);
          v.x += t.x;
          v.y += t.y;
          v.z += t.z;
          v.w += t.w;
        }
        const int u = u0 + idx;
        const int m = u / (NTILE / 4), n = (u % (NTILE / 4)) * 4;
        alignas(8) scalar_t o[4] = {ScalarTraits<scalar_t>::from_float(v.x), ScalarTraits<scalar_t>::from_float(v.y),
                                    ScalarTraits<scalar_t>::from_float(v.z), ScalarTraits<scalar_t>::from_float(v.w)};
        *reinterpret_cast<uint2*>(crow + (int64_t)m * p.N + n) = *reinterpret_cast<const uint2*>(o);
      }
      named_bar_sync(3, MT);  // the chunk buffer is reused
    }
    if (mtid == 0) p.counters[tile] = 0;  // leave the caller's lock buffer all-zero
  };
  auto merge_phase = [&](int mtid) {
    named_bar_sync(3, MT);  // every segment is done: the rings are idle and the list of tiles to merge is published
    const int n = s_pend[0];
    for (int i = 0; i < n; ++i) merge_tile(s_pend[1 + i], mtid);
  };

#ifdef APH_GEMM_TRACE
  if (p.trace != nullptr && blockIdx.x == 0 && tid == 0) p.trace[512 + 48 + 2] = clock64();
#endif
  if (w == W_PROD) {
    // ==================== weight producer: packed weights, scales, zero points ====================
    if (lane != 0) {
      if (CS) cluster_sync_twice_unaligned();
      return;
    }
    prefetch_tensormap(&tmB);
    if (grouped) prefetch_tensormap(&tmS);
    if (HAS_ZP && grouped) prefetch_tensormap(&tmZ);
    const uint64_t pol_stream = policy_evict_first();  // weights are read once (or re-read from L2 by the next batch tile)
    const uint32_t stage_tx = (uint32_t)(W_BYTES + gps * NTILE * 2 + (HAS_ZP ? (ZPF ? gps * NTILE * 2 : gps * (NTILE / PACK) * 4) : 0));
    int it = 0;
    RingCursor<SW> ring;  // the producer handles every stage
    auto issue = [&](int sw, int slice, int ks) {
      const uint32_t dst = smem_base + C::W_OFF + sw * W_STAGE;
      const uint32_t fb = fullw_bar + sw * 8;
      mbar_expect_tx(fb, stage_tx);
      tma_load_2d_hint(dst, &tmB, slice * NTILE * (BITS / 4), ks * 8, fb, pol_stream);  // 8-byte words: 1 (2) per column at 4 (8) bit
      if (grouped) {
        tma_load_2d(dst + W_BYTES, &tmS, slice * NTILE, ks * gps, fb);
        if constexpr (HAS_ZP) tma_load_2d(dst + W_BYTES + S_BYTES, &tmZ, ZPF ? slice * NTILE : slice * (NTILE / PACK), ks * gps, fb);
      }
    };
#ifdef APH_RING_FAULT
    int held_sw = -1, held_slice = 0, held_ks = 0;  // an odd stage held back while the even stage behind it goes out first
#endif
    APH_SEGMENTS {
      APH_SEG_VARS
#pragma unroll 1
      for (int ks = k0; ks < k1; ++ks, ++it) {
        const int sw = ring.slot(it);
        mbar_wait(emptyw_bar + sw * 8, ring.phase(it) ^ 1);
        ring.advance(it);
#ifdef APH_RING_FAULT
        if (it & 1) {
          held_sw = sw;
          held_slice = slice;
          held_ks = ks;
          continue;
        }
#endif
        issue(sw, slice, ks);
#ifdef APH_RING_FAULT
        if (held_sw >= 0) {
          ring_fault_delay();
          issue(held_sw, held_slice, held_ks);
          held_sw = -1;
        }
#endif
      }
    }
#ifdef APH_RING_FAULT
    if (held_sw >= 0) issue(held_sw, held_slice, held_ks);
#endif
    if (CS) cluster_sync_twice_unaligned();
    return;
  }

  if (w == W_ACT) {
    // ==================== activation producer ====================
    if (lane != 0) {
      if (CS) cluster_sync_twice_unaligned();
      return;
    }
    prefetch_tensormap(&tmA);
    griddep_wait();  // activations are the previous kernel's output
    const uint64_t pol_keep = policy_evict_last();  // activations are shared by every CTA
    int it = 0;
    RingCursor<SA> ring;  // handles every stage
    APH_SEGMENTS {
      APH_SEG_VARS
#pragma unroll 1
      for (int ks = k0; ks < k1; ++ks, ++it) {
        const int sa = ring.slot(it);
        mbar_wait(emptya_bar + sa * 8, ring.phase(it) ^ 1);
        ring.advance(it);
        const uint32_t dst = smem_base + C::ACT_OFF + sa * C::ACT_BYTES;
        const uint32_t fb = ready_bar + sa * 8;
        mbar_expect_tx(fb, (uint32_t)C::ACT_BYTES);
        tma_load_2d_hint(dst, &tmA, ks * KSTAGE, mt * MB, fb, pol_keep);
        tma_load_2d_hint(dst + C::ACT_BOX, &tmA, ks * KSTAGE + 64, mt * MB, fb, pol_keep);
      }
    }
    if (CS) cluster_sync_twice_unaligned();
    return;
  }

  if (w == W_MMA || (C::NI == 2 && w == W_MMA2)) {
    // ============================ MMA issuer(s): one elected lane of a convergent warp ============================
    {
      constexpr int NI = C::NI;
      const int mi = (w == W_MMA) ? 0 : 1;  // takes the stages with it % NI == mi
      constexpr uint32_t DESC_HI = (uint32_t)(1024 >> 4) | (1u << 14) | (2u << 29);  // SBO, version 1, SWIZZLE_128B
      // This CTA owns all 512 TMEM columns, so its allocation starts at lane 0, column 0.
      if (tmem_base != 0) __trap();
      constexpr uint32_t tmem0 = 0;
      int it = 0;  // global index of the segment's first stage
      bool ok = false;  // result of an early, non-blocking look at the barrier of this issuer's next stage
      // NI == 1: every stage; NI == 2: the stages of group mi, i.e. those with (index & 1) == mi
      using Cursor = typename std::conditional<NI == 2, GroupCursor<SA>, RingCursor<SA>>::type;
      Cursor ring(mi);
      APH_SEGMENTS {
        APH_SEG_VARS
        const int db = seg % NDBUF;
        mbar_wait(dfree_bar + db * 8, ((seg / NDBUF) & 1) ^ 1);  // the epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem0 + (db * NI + mi) * MB;
        uint32_t acc = 0;
        const int it_seg = it;
        it += k1 - k0;
#pragma unroll 1
        for (int ks = k0 + (NI == 2 ? ((mi ^ it_seg) & 1) : 0); ks < k1; ks += NI) {
          const int its = it_seg + (ks - k0);
          const int sa = ring.slot(its);
#ifdef APH_GEMM_TRACE
          const bool tr = p.trace != nullptr && blockIdx.x == 0 && its < 64 && lane == 0;
          if (tr) p.trace[its * 8 + 0] = clock64();
#endif
          if (!ok) mbar_wait(ready_bar + sa * 8, ring.phase(its));  // activations landed and dequantised weights are in TMEM
          ring.advance(its);
          tc_fence_after();
#ifdef APH_GEMM_TRACE
          if (tr) p.trace[its * 8 + 2] = clock64();
#endif
          const uint32_t desc_lo = (((smem_base + C::ACT_OFF + sa * C::ACT_BYTES) >> 4) & 0x3fffu) | (1u << 16);
          const uint32_t a_tmem = tmem0 + C::A_COL0 + sa * 64;
          const int nx = its + NI;
          ok = mbar_test(ready_bar + ring.slot(nx) * 8, ring.phase(nx));  // poll the next stage while these MMAs are issued
          tc_mma_ts_stage8(d_tmem, a_tmem, desc_lo, DESC_HI, C::IDESC, acc, (uint32_t)(C::ACT_BOX >> 4));
          acc = 1;
          tc_commit_elect(emptya_bar + sa * 8);  // slot (activations + TMEM weights) free once these MMAs retire
#ifdef APH_GEMM_TRACE
          if (tr) p.trace[its * 8 + 3] = clock64();
#endif
        }
        tc_commit_elect(done_bar + db * 8);
      }
    }
    if (w != W_MMA) {
      if (CS) cluster_sync_twice_unaligned();
      return;
    }
    __syncwarp();
    if (CS) {
      // every thread of the cluster takes part in the two phases of the epilogue's merge (see there)
      cluster_arrive_release();
      cluster_wait_acquire();
      cluster_arrive_release();
      cluster_wait_acquire();
    }
    named_bar_sync(2, (EPI_WARPS + 1) * 32);  // every TMEM read of the epilogue is done
    tmem_dealloc(tmem_base, C::TMEM_COLS);
    return;
  }

  if (w >= W_DQ0) {
    // ================================ dequantiser warps ================================
    griddep_wait();                       // they join the end-of-kernel merge (global reads / writes)
    const int qd = w & 3;                 // TMEM lane quadrant this warp may access
    const int dw = w - W_DQ0;
    const int h = (dw >> 2) & 1;          // which half of the stage's k16-rows
    const int grp = dw >> 3;              // which stages: it & 1 == grp
    const int g = lane >> 2;
    // this lane's words of a k16-row: 4 bit - one int32 per 16-column tile, the quadrant's two tiles are 8 bytes;
    // 8 bit - two int32 per tile (gptq_marlin.cu:274-368), the quadrant's two tiles are 16 bytes
    const uint32_t offW = (uint32_t)(C::W_OFF + (4 * h) * ROW_BYTES + (qd >> 1) * (ROW_BYTES / 2) + lane * (4 * BITS) + (qd & 1) * (2 * BITS));
    const uint32_t offS = (uint32_t)(C::W_OFF + W_BYTES + (qd >> 1) * 128 + g * 16 + (qd & 1) * 8);
    const uint32_t offZ = BITS == 4 ? (uint32_t)(C::W_OFF + W_BYTES + S_BYTES + (qd >> 1) * 32 + g * 4)
                                    : (uint32_t)(C::W_OFF + W_BYTES + S_BYTES + (qd >> 1) * 64 + g * 8 + (qd & 1) * 4);
    const uint32_t a_lane = (uint32_t)(32 * qd) << 16;
    const int rpg_shift = G128 ? 3 : grouped ? 31 - __clz(p.group_size / 16) : 3;  // log2(k16-rows per group): 1, 2, 3
    const bool two_groups = !G128 && grouped && p.group_size == 32;                // a warp's 4 rows span two groups

    // what is subtracted from the spliced value, per 16-row tile for rows g and g + 8. 16-bit route (4 bit; 8 bit f16):
    // packed pair (MAGIC + z, MAGIC + z). 8-bit bf16 goes through fp32 (2^23 + byte): the fp32 constant 2^23 + z.
    constexpr bool VIA_F32 = BITS == 8 && !F16;
    constexpr uint32_t BIAS = BITS == 4 ? 0x00080008u : 0x00800080u;  // the symmetric types: uint4b8, uint8b128
    uint32_t s_lo[2], s_hi[2], c_lo[2], c_hi[2];
    uint32_t z_lo[2] = {0u, 0u}, z_hi[2] = {0u, 0u};  // ZPF only
    (void)z_lo;
    (void)z_hi;
    c_lo[0] = c_lo[1] = c_hi[0] = c_hi[1] = ZPF ? C::MAGIC : VIA_F32 ? 0x4B000080u : (C::MAGIC | BIAS);
    s_lo[0] = s_lo[1] = s_hi[0] = s_hi[1] = 0;
    auto set_zp = [&](uint32_t zword) {
#pragma unroll
      for (int tt = 0; tt < 2; ++tt) {
        uint32_t zl, zh;
        if constexpr (BITS == 4) {
          // nibble e' of the packed word holds column block {0,2,4,6,1,3,5,7}[e']: tile j rows g / g+8 sit in
          // nibbles j and j + 4
          const int j = 2 * (qd & 1) + tt;
          zl = (zword >> (4 * j)) & 0xf;
          zh = (zword >> (4 * j + 16)) & 0xf;
        } else {
          // 8 bit: four zero points per word, interleave {0,2,1,3}: this quadrant's word holds its two tiles, tile tt
          // rows g / g+8 in bytes tt and tt + 2
          zl = (zword >> (8 * tt)) & 0xff;
          zh = (zword >> (8 * tt + 16)) & 0xff;
        }
        if constexpr (VIA_F32) {
          c_lo[tt] = 0x4B000000u | zl;
          c_hi[tt] = 0x4B000000u | zh;
        } else {
          c_lo[tt] = C::MAGIC | zl | (zl << 16);
          c_hi[tt] = C::MAGIC | zh | (zh << 16);
        }
      }
    };
    auto load_group = [&](uint32_t wstage, int gl) {
      uint32_t sw0, sw1;
      lds64(wstage + offS + gl * (NTILE * 2), sw0, sw1);
      s_lo[0] = __byte_perm(sw0, sw0, 0x1010);
      s_hi[0] = __byte_perm(sw0, sw0, 0x3232);
      s_lo[1] = __byte_perm(sw1, sw1, 0x1010);
      s_hi[1] = __byte_perm(sw1, sw1, 0x3232);
      if constexpr (ZPF) {
        // float zero points: same layout as the scales; the subtraction takes two exact / rounded steps (q, then q - z)
        lds64(wstage + offS + S_BYTES + gl * (NTILE * 2), sw0, sw1);
        z_lo[0] = __byte_perm(sw0, sw0, 0x1010);
        z_hi[0] = __byte_perm(sw0, sw0, 0x3232);
        z_lo[1] = __byte_perm(sw1, sw1, 0x1010);
        z_hi[1] = __byte_perm(sw1, sw1, 0x3232);
      } else if constexpr (HAS_ZP) {
        set_zp(lds32(wstage + offZ + gl * (NTILE / PACK) * 4));
      }
    };
    // two bytes of an 8-bit word (b, b + 2) -> the exact 16-bit pair (q - z) [* s]
    auto pair8 = [&](uint32_t word, int b, uint32_t cc, uint32_t ss) -> uint32_t {
      if constexpr (VIA_F32) {
        const float f0 = __uint_as_float(__byte_perm(word, 0x4B000000u, b == 0 ? 0x7650 : 0x7651)) - __uint_as_float(cc);
        const float f1 = __uint_as_float(__byte_perm(word, 0x4B000000u, b == 0 ? 0x7652 : 0x7653)) - __uint_as_float(cc);
        uint32_t v = __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632);  // small integers: the top halves are exact bf16
        if (grouped) {
          __nv_bfloat162 t = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&v), *reinterpret_cast<__nv_bfloat162*>(&ss));
          v = *reinterpret_cast<uint32_t*>(&t);
        }
        return v;
      } else {
        return sub_mul2<F16>(__byte_perm(word, C::MAGIC, b == 0 ? 0x7250 : 0x7351), cc, ss, grouped);
      }
    };

    int pending = -1;  // activation/TMEM slot whose stores are issued but not yet published to the MMA warp
    auto publish = [&]() {
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(ready_bar + pending * 8);
    };

    int it = 0;  // global index of the segment's first stage
    GroupCursor<SW> ring_w(grp);  // this warp handles the stages of group grp only: the stages with (index & 1) == grp
    GroupCursor<SA> ring_a(grp);
    APH_SEGMENTS {
      APH_SEG_VARS
      if (HAS_ZP && !ZPF && !grouped)
        set_zp(*reinterpret_cast<const uint32_t*>(
            p.zp + (BITS == 4 ? ((int64_t)slice * (NTILE / 8) + (qd >> 1) * 8 + g) * 4
                              : ((int64_t)slice * (NTILE / 4) + (qd >> 1) * 16 + g * 2 + (qd & 1)) * 4)));
      const int it_seg = it;
      it += k1 - k0;
#pragma unroll 1
      for (int ks = k0 + ((grp ^ it_seg) & 1); ks < k1; ks += 2) {
        const int its = it_seg + (ks - k0);  // (its & 1) == grp
        (void)its;
        const int sw = ring_w.slot(its), sa = ring_a.slot(its);
        const uint32_t ph_w = ring_w.phase(its), ph_a = ring_a.phase(its);
        ring_w.advance();
        ring_a.advance();
#ifdef APH_GEMM_TRACE
        const bool tr = p.trace != nullptr && blockIdx.x == 0 && its < 64 && (dw & 7) == 0 && lane == 0;
        if (tr) p.trace[its * 8 + 4] = clock64();
#endif
        mbar_wait(fullw_bar + sw * 8, ph_w);
#ifdef APH_GEMM_TRACE
        if (tr) p.trace[its * 8 + 5] = clock64();
#endif
        const uint32_t wstage = smem_base + sw * W_STAGE;
        uint32_t q[4][BITS / 2];  // 4 bit: one word per tile; 8 bit: two words per tile
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if constexpr (BITS == 4)
            lds64(wstage + offW + i * ROW_BYTES, q[i][0], q[i][1]);
          else
            lds128(wstage + offW + i * ROW_BYTES, q[i][0], q[i][1], q[i][2], q[i][3]);
        }
        // the previous stage's TMEM stores had the wait and the loads above to land: hand that slot to the MMA warps
        // before starting on this stage's arithmetic (publishing later would keep one more slot per group in limbo)
        if (pending >= 0) publish();
        uint32_t a[2][16];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if constexpr (G128) {
            if (i == 0) load_group(wstage, 0);
          } else {
            if (grouped && (i == 0 || (i == 2 && two_groups))) load_group(wstage, (4 * h + i) >> rpg_shift);
          }
#pragma unroll
          for (int tt = 0; tt < 2; ++tt) {
            if constexpr (BITS == 4) {
              const uint32_t a0 = and_or(q[i][tt], 0x000f000fu, C::MAGIC);
              const uint32_t a2 = and_or(q[i][tt] >> 4, 0x000f000fu, C::MAGIC);
              const uint32_t a1 = and_or(q[i][tt] >> 8, 0x000f000fu, C::MAGIC);
              const uint32_t a3 = and_or(q[i][tt] >> 12, 0x000f000fu, C::MAGIC);
              if constexpr (ZPF) {
                a[tt][4 * i + 0] = sub_mul2<F16>(sub_mul2<F16>(a0, c_lo[tt], 0u, false), z_lo[tt], s_lo[tt], true);
                a[tt][4 * i + 1] = sub_mul2<F16>(sub_mul2<F16>(a1, c_hi[tt], 0u, false), z_hi[tt], s_hi[tt], true);
                a[tt][4 * i + 2] = sub_mul2<F16>(sub_mul2<F16>(a2, c_lo[tt], 0u, false), z_lo[tt], s_lo[tt], true);
                a[tt][4 * i + 3] = sub_mul2<F16>(sub_mul2<F16>(a3, c_hi[tt], 0u, false), z_hi[tt], s_hi[tt], true);
              } else {
                a[tt][4 * i + 0] = sub_mul2<F16>(a0, c_lo[tt], s_lo[tt], G128 || grouped);
                a[tt][4 * i + 1] = sub_mul2<F16>(a1, c_hi[tt], s_hi[tt], G128 || grouped);
                a[tt][4 * i + 2] = sub_mul2<F16>(a2, c_lo[tt], s_lo[tt], G128 || grouped);
                a[tt][4 * i + 3] = sub_mul2<F16>(a3, c_hi[tt], s_hi[tt], G128 || grouped);
              }
            } else {
              // word 0 of a tile: row g, bytes (0, 2) = k low pair, (1, 3) = k high pair; word 1: row g + 8
              a[tt][4 * i + 0] = pair8(q[i][2 * tt], 0, c_lo[tt], s_lo[tt]);
              a[tt][4 * i + 2] = pair8(q[i][2 * tt], 1, c_lo[tt], s_lo[tt]);
              a[tt][4 * i + 1] = pair8(q[i][2 * tt + 1], 0, c_hi[tt], s_hi[tt]);
              a[tt][4 * i + 3] = pair8(q[i][2 * tt + 1], 1, c_hi[tt], s_hi[tt]);
            }
          }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(emptyw_bar + sw * 8);  // this warp's words of the stage are in registers
#ifdef APH_GEMM_TRACE
        if (tr) p.trace[its * 8 + 6] = clock64();
#endif
        mbar_wait(emptya_bar + sa * 8, ph_a ^ 1);  // the MMAs that read this TMEM slot have retired
        tc_fence_after();
#ifdef APH_GEMM_TRACE
        if (tr) p.trace[its * 8 + 7] = clock64();
#endif
        const uint32_t a_tmem = tmem_base + a_lane + C::A_COL0 + sa * 64 + (4 * h) * 8;
        tmem_st_16x128b_x8(a_tmem, a[0]);
        tmem_st_16x128b_x8(a_tmem + (16u << 16), a[1]);
        pending = sa;
      }
    }
    if (pending >= 0) publish();
    if (CS) {  // cluster split-K: the epilogue warps merge through shared memory, nothing to do here but to be counted
      cluster_sync_twice_unaligned();
      return;
    }
    merge_phase(EPI_WARPS * 32 + tid - W_DQ0 * 32);
    return;
  }
  if (w < W_EPI0) {  // second issuer warp idles when NI == 1
    if (CS) cluster_sync_twice_unaligned();
    return;
  }

  // ================================ epilogue warps ================================
  {
    griddep_wait();  // global writes (C, partials, counters) only after the previous kernel is complete
    const int qd = w & 3;
    const int etid = tid - W_EPI0 * 32;
    constexpr int ET = EPI_WARPS * 32;
    const uint32_t d_lane = (uint32_t)(32 * qd) << 16;
    const int n_local = 32 * qd + lane;  // TMEM lane == output column inside the slice
    const int b = blockIdx.x;
    uint8_t* stage = smem + C::STAGE_OFF;
    int pend[2], npend = 0;  // shared tiles this CTA was the last to contribute to (at most its first and last)
    int it0 = 0;             // global stage index of the segment's first stage
    APH_SEGMENTS {
      APH_SEG_VARS
      const int db = seg % NDBUF;
      // the second issuer's accumulator holds this segment's odd (or even) stages - if it got any
      const bool two = C::NI == 2 && (k1 - k0) > 1;
      const int lone = it0 % C::NI;  // with a single stage: the issuer that took it
      it0 += k1 - k0;
      const bool direct = (k0 == 0 && k1 == KS);
      const int m_rows = min(MB, p.M - mt * MB);
      float s_n = 1.f;
      if (!grouped) {
        // channelwise scales use the "single" permutation over 32-column blocks (marlin_utils.py:204-213)
        const int nn = n_local & 63, j = nn >> 4, rr = nn & 15, gg = rr & 7;
        const int pos = (j >> 1) * 32 + 8 * (gg >> 1) + (gg & 1) + 4 * (j & 1) + ((rr >> 3) << 1);
        s_n = ScalarTraits<scalar_t>::to_float(reinterpret_cast<const scalar_t*>(p.s)[(int64_t)slice * NTILE + (n_local & 64) + pos]);
      }
      scalar_t* crow = reinterpret_cast<scalar_t*>(p.c) + (int64_t)mt * MB * p.N + (int64_t)slice * NTILE;
      float* mine = p.scratch + ((size_t)b * 2 + (segj == 0 ? 0 : 1)) * (size_t)(MB * NTILE) + n_local;

#ifdef APH_GEMM_TRACE
      const bool etr = p.trace != nullptr && b == 0 && etid == 0 && seg < 6;
#endif
#ifdef APH_GEMM_TRACE
      if (etr) p.trace[512 + seg * 8 + 0] = clock64();
#endif
      // The accumulator is complete one main-loop pass over this segment from now (~350 ns per stage). A waiting warp
      // is woken by every barrier event of the CTA, and four warps polling through the whole pass took 7 % of the
      // kernel's issue slots from the dequantisers (profiles/r02_ncu_gemm_tc_gate_up_m64_g128.txt): sleep through most
      // of the pass first (__nanosleep(t) returns within 2 t: 100 ns per stage asks for at most 200).
      if (!mbar_test(done_bar + db * 8, (seg / NDBUF) & 1)) {
        __nanosleep((unsigned)min((k1 - k0) * 100, 200000));
        mbar_wait(done_bar + db * 8, (seg / NDBUF) & 1);
      }
      tc_fence_after();
#ifdef APH_GEMM_TRACE
      if (etr) p.trace[512 + seg * 8 + 1] = clock64();
#endif
      const uint32_t d0 = tmem_base + d_lane + (db * C::NI + (two ? 0 : lone)) * MB;
      const uint32_t d1 = tmem_base + d_lane + (db * C::NI + 1) * MB;
      if (CS) {
        // ---- cluster split-K: the CS CTAs of the cluster hold partial sums of the same tile. Chunk cc (16 batch rows)
        //      belongs to rank cc % CS: everybody pushes its other chunks into the owner's (now idle) pipeline rings
        //      through DSMEM, the owner adds the CS partials in rank order and writes C. No scratch, no atomics.
        //      Barrier phase one = "my MMAs are complete, my rings are idle" (also: every CTA of the cluster runs). ----
        const int crank = (int)cluster_ctarank();
        const int chunks = (m_rows + 15) / 16;
        const int owned_max = (MB / 16 + CS - 1) / CS;
        float* recv = reinterpret_cast<float*>(smem);  // [source rank][owned chunk][128 columns][16 rows]
        auto load_chunk = [&](int cc, uint32_t (&v)[16]) {
          tmem_ld_32x32b_x16(d0 + 16 * cc, v);
          if (two) {
            uint32_t v2[16];
            tmem_ld_32x32b_x16(d1 + 16 * cc, v2);
            tmem_wait_ld();
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]));
          } else {
            tmem_wait_ld();
          }
        };
        cluster_arrive_release();
        cluster_wait_acquire();
        const uint32_t recv_u32 = smem_base;
#pragma unroll 1
        for (int cc = 0; cc < chunks; ++cc) {
          const int owner = cc % CS;
          if (owner == crank) continue;
          uint32_t v[16];
          load_chunk(cc, v);
          const uint32_t dst = cluster_map_shared(recv_u32 + (uint32_t)(((crank * owned_max + cc / CS) * NTILE + n_local) * 64), (uint32_t)owner);
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + j * 4), "r"(v[j]), "r"(v[j + 1]), "r"(v[j + 2]),
                         "r"(v[j + 3])
                         : "memory");
        }
        cluster_arrive_release();
        cluster_wait_acquire();
#pragma unroll 1
        for (int cc = crank; cc < chunks; cc += CS) {
          const int rows = m_rows - 16 * cc;
          uint32_t v[16];
          load_chunk(cc, v);
          float acc[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = 0.f;
#pragma unroll 1
          for (int r = 0; r < CS; ++r) {  // rank order: deterministic
            if (r == crank) {
#pragma unroll
              for (int j = 0; j < 16; ++j) acc[j] += __uint_as_float(v[j]);
            } else {
              const float4* src = reinterpret_cast<const float4*>(recv + (size_t)((r * owned_max + cc / CS) * NTILE + n_local) * 16);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float4 t = src[j];
                acc[4 * j] += t.x;
                acc[4 * j + 1] += t.y;
                acc[4 * j + 2] += t.z;
                acc[4 * j + 3] += t.w;
              }
            }
          }
#pragma unroll
          for (int j = 0; j < 16; ++j)
            *reinterpret_cast<scalar_t*>(stage + j * (NTILE * 2) + n_local * 2) = ScalarTraits<scalar_t>::from_float(acc[j] * s_n);
          named_bar_sync(1, ET);
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int vec = etid + i * ET, row = vec >> 4, c16 = vec & 15;
            if (row < rows)
              *reinterpret_cast<uint4*>(crow + (int64_t)(16 * cc + row) * p.N + c16 * 8) =
                  *reinterpret_cast<const uint4*>(stage + row * (NTILE * 2) + c16 * 16);
          }
          named_bar_sync(1, ET);  // the staging tile is reused by the next chunk
        }
        tc_fence_before();
        continue;  // exactly one segment per CTA in this mode
      }
#pragma unroll 1
      for (int cc = 0; cc < MB / 16; ++cc) {
        const int rows = m_rows - 16 * cc;  // batch rows in this chunk of 16 accumulator columns
        if (rows <= 0) break;
        uint32_t v[16];
        tmem_ld_32x32b_x16(d0 + 16 * cc, v);
        if (two) {
          uint32_t v2[16];
          tmem_ld_32x32b_x16(d1 + 16 * cc, v2);
          tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]));
        } else {
          tmem_wait_ld();
        }
        if (cc == MB / 16 - 1 || rows <= 16) {
          // last TMEM read of this accumulator: the MMA warps may start the next tile in it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(dfree_bar + db * 8);
        }
        if (direct) {
          // transpose through shared memory so that C rows leave as 16-byte vectors
#pragma unroll
          for (int j = 0; j < 16; ++j)
            *reinterpret_cast<scalar_t*>(stage + j * (NTILE * 2) + n_local * 2) = ScalarTraits<scalar_t>::from_float(__uint_as_float(v[j]) * s_n);
          named_bar_sync(1, ET);
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int vec = etid + i * ET, row = vec >> 4, c16 = vec & 15;
            if (row < rows)
              *reinterpret_cast<uint4*>(crow + (int64_t)(16 * cc + row) * p.N + c16 * 8) =
                  *reinterpret_cast<const uint4*>(stage + row * (NTILE * 2) + c16 * 16);
          }
          named_bar_sync(1, ET);  // the staging tile is reused by the next chunk
        } else {
          float* dst = mine + (size_t)(16 * cc) * NTILE;
          if (rows >= 16) {
#pragma unroll
            for (int j = 0; j < 16; ++j) __stcg(dst + j * NTILE, __uint_as_float(v[j]) * s_n);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (j < rows) __stcg(dst + j * NTILE, __uint_as_float(v[j]) * s_n);
          }
        }
      }
#ifdef APH_GEMM_TRACE
      if (etr) p.trace[512 + seg * 8 + 2] = clock64();
#endif
      if (direct) continue;

      // ---- the tile's k range is shared with other CTAs: whoever arrives last merges the partials (at the end) ----
      __threadfence();
      named_bar_sync(1, ET);
      if (etid == 0) {
        const int n_c = owner_of(tile * KS + KS - 1, T, G) - owner_of(tile * KS, T, G) + 1;
        *s_last = (atomicAdd(&p.counters[tile], 1) == n_c - 1);
      }
      named_bar_sync(1, ET);
      if (*s_last) pend[npend++] = tile;
      named_bar_sync(1, ET);  // s_last may be rewritten by the next segment
#ifdef APH_GEMM_TRACE
      if (etr) p.trace[512 + seg * 8 + 3] = clock64();
#endif
    }
    if (etid == 0) {
      s_pend[0] = npend;
      s_pend[1] = pend[0];
      s_pend[2] = pend[1];
    }
#ifdef APH_GEMM_TRACE
    if (p.trace != nullptr && b == 0 && etid == 0) p.trace[512 + 48 + 0] = clock64();
#endif
    if (!CS) merge_phase(etid);
#ifdef APH_GEMM_TRACE
    if (p.trace != nullptr && b == 0 && etid == 0) p.trace[512 + 48 + 1] = clock64();
#endif
    tc_fence_before();
    named_bar_arrive(2, (EPI_WARPS + 1) * 32);
  }
#undef APH_SEGMENTS
#undef APH_SEG_VARS
}

}  // namespace mgtc
}  // namespace aphb200
