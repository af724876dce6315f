// W4A16 dequant-GEMM kernel on Marlin-format weights (small / medium batch path, mma.sync).
// See marlin_gemm.cu for the design summary.
#pragma once
#include <type_traits>

#include "common.cuh"
#include "ptx.cuh"

namespace aphb200 {
namespace mg {

using namespace ::aphb200::ptx;

constexpr int KT = 128;        // k rows per warp block (= 8 k16-rows of the Marlin layout)
constexpr int CONSUMERS = 4;   // consumer warps working on one stage
constexpr int SG = 2;          // stage groups: the warps of group sg consume stages it % SG == sg (latency hiding)
constexpr int THREADS = (SG * CONSUMERS + 1) * 32;  // + 1 producer warp

struct Params {
  const uint8_t* s;   // scales, bytes (channelwise path reads them directly)
  const uint8_t* zp;  // zero points, bytes (or null)
  void* c;
  float* scratch;      // [grid][2][MP*NT] fp32 partials: slot 0 = head of a CTA's item range, slot 1 = its tail
  int* counters;       // [n_slices] arrival counters, all-zero on entry and on exit
  int N, K;
  int m0, m_rows;      // this launch covers rows [m0, m0 + m_rows) of a / c
  int group_size;      // == K for channelwise
  int k_stages;        // ceil(K / (128 * KG)) pipeline stages over the whole K
  int total;           // n_slices * k_stages work items (slice-major); CTA i owns [i*total/grid, (i+1)*total/grid)
  int dbg;             // development switch (APH_GEMM_DBG): 1 = skip the math, 2 = skip the loads
  unsigned long long* trace;  // development (APH_GEMM_TRACE builds): [grid][8] globaltimer stamps, else null
  int cluster_split;   // S >= 2: aligned split-K plan launched as thread-block clusters of S CTAs (the S pieces of one
                       // slice); their partials meet in rank 0's shared memory instead of the scratch. 0: scratch + counters
};

// One pipeline stage holds KG x NWN warp blocks of (128 k x 64 n): 16 KiB of weights whatever the slice
// width, fetched by ONE tiled TMA instruction; every consumer warp owns one block of every stage.
template <typename scalar_t, int NB, int NT>
struct Cfg {
  static constexpr bool F16 = std::is_same<scalar_t, __half>::value;
  static constexpr int NWN = NT / 64;                 // consumer warps along n (64 columns each)
  static constexpr int KG = CONSUMERS / NWN;          // consumer warps along k (128 k each)
  static constexpr int KS = KT * KG;                  // k extent of a stage
  static constexpr int MP = 8 * NB;                   // padded batch rows per launch
  static constexpr int B_BYTES = 8 * KG * NT * 8;     // [8*KG k16-rows][NT columns] x 8 bytes = 16 KiB
  static constexpr int A_BOX = MP * 128;              // one activation box: MP rows x 64 k, 128B-swizzled
  static constexpr int A_BYTES = 2 * KG * A_BOX;
  static constexpr int S_BYTES = 4 * KG * NT * 2;     // up to 4 groups per 128 k
  static constexpr int Z_BYTES = 4 * KG * (NT / 8) * 4;
  static constexpr int STAGE_BYTES = (B_BYTES + A_BYTES + S_BYTES + Z_BYTES + 1023) / 1024 * 1024;
  static constexpr int RP = NT + 4;                   // padded row of the reduction buffer (floats)
  static constexpr int RED_BYTES = SG * KG * MP * RP * 4;
  static constexpr int CTAS_PER_SM = NB <= 2 ? 2 : 1;
  static constexpr int BUDGET = (224 * 1024) / CTAS_PER_SM - 2048 - RED_BYTES - 256;
  static constexpr int STAGES = BUDGET / STAGE_BYTES > 8 ? 8 : BUDGET / STAGE_BYTES;
  static_assert(STAGES >= 2, "pipeline too shallow");
  static constexpr int BAR_BYTES = 2 * STAGES * 8 + 64;  // barriers + fix-up flags of the (at most two) shared slices
  static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + RED_BYTES + BAR_BYTES;
  static constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;  // 1024 / 128 as packed pairs
};

// The pieces ("segments") of output slices inside the item range [b, e) of one CTA, in PROCESSING order: the
// pieces that share their slice with a neighbouring CTA first (at most two: the head and the tail of the range),
// whole slices afterwards. The arrival on a shared slice is thus published early and the last arriver, who sums the
// partials, finds them complete when it gets there at the end of its own whole slices: the fix-up round trip
// (fence + atomic + reload) hides behind streaming instead of ending every CTA. f(slice, k0, cnt, whole).
template <typename F>
__device__ __forceinline__ void for_each_segment(int b, int e, int ks, F&& f) {
  const int s_first = b / ks, s_last = (e - 1) / ks;
#pragma unroll 1
  for (int pass = 0; pass < 2; ++pass) {
#pragma unroll 1
    for (int sl = s_first; sl <= s_last; ++sl) {
      const int sb = max(b, sl * ks), se = min(e, (sl + 1) * ks);
      const bool whole = (se - sb) == ks;
      if (whole == (pass == 1)) f(sl, sb - sl * ks, se - sb, whole);
    }
  }
}

// Deferred fix-up of the (at most two) shared slices of a CTA, run by all consumer threads after their last segment.
// Kept out of line so that its registers do not count against the main loop's budget (96 per thread at two CTAs of
// nine warps per SM).
template <typename scalar_t, int MP, int NT, int CT>
__device__ __noinline__ void sum_shared_slices(scalar_t* c, const float* scratch, int* counters, int N, int m_rows, int k_stages,
                                               int total, const int* s_flag, int n_shared, int* list, int tid) {
#pragma unroll 1
  for (int j = 0; j < n_shared; ++j) {
    if (!s_flag[4 * j + 3]) continue;  // CTA-uniform: somebody else arrived last
    const int slice = s_flag[4 * j], first = s_flag[4 * j + 1], n_parts = s_flag[4 * j + 2];
    const int s_begin = slice * k_stages;
    // scratch tile index of every contributor
    for (int ci = tid; ci < n_parts; ci += CT) {
      const uint32_t cta = (uint32_t)(first + ci);
      const int cta_begin = (int)(cta * (uint32_t)total / gridDim.x);
      list[ci] = (int)cta * 2 + (cta_begin < s_begin ? 1 : 0);
    }
    fence_acq_rel_gpu();
    named_bar_sync(1, CT);
    const int nelem = m_rows * NT;
    // up to 4 elements x 4 contributors = 16 independent loads in flight per thread (L2 latency, not bandwidth, is the cost)
    for (int idx0 = tid; idx0 < nelem; idx0 += 4 * CT) {
      float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
      for (int c0 = 0; c0 < n_parts; c0 += 4) {
        float t[4][4];
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const float* src = scratch + (size_t)list[min(c0 + cc, n_parts - 1)] * (size_t)(MP * NT);
#pragma unroll
          for (int u = 0; u < 4; ++u) t[cc][u] = (c0 + cc < n_parts && idx0 + u * CT < nelem) ? __ldcg(src + idx0 + u * CT) : 0.f;
        }
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)  // fixed order: deterministic sums
#pragma unroll
          for (int u = 0; u < 4; ++u) v[u] += t[cc][u];
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int idx = idx0 + u * CT;
        if (idx < nelem) c[(int64_t)slice * NT + (int64_t)(idx / NT) * N + idx % NT] = ScalarTraits<scalar_t>::from_float(v[u]);
      }
    }
    if (tid == 0) counters[slice] = 0;  // leave the caller's lock buffer all-zero
    named_bar_sync(1, CT);                // `list` is reused
  }
}

// Tensor maps (built on the host per call, marlin_gemm.cu):
//   tmB: weights as a 2-D array of 8-byte words [K/16][N], box [8*KG][NT]
//   tmA: activations [M][K] 16-bit, box [MP][64], SWIZZLE_128B, rows >= M read as zero
//   tmS: scales [groups][N] 16-bit, box [KG*gps][NT];  tmZ: zero points [groups][N/8] 32-bit, box [KG*gps][NT/8]
#ifdef APH_GEMM_TRACE
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define APH_MMA_STAMP(slot, cond)                                              \
  do {                                                                         \
    if (p.trace != nullptr && (cond)) p.trace[blockIdx.x * 8 + (slot)] = global_ns(); \
  } while (0)
#else
#define APH_MMA_STAMP(slot, cond) \
  do {                            \
  } while (0)
#endif

template <typename scalar_t, int NB, int NT, bool HAS_ZP>
__global__ void __launch_bounds__(THREADS, Cfg<scalar_t, NB, NT>::CTAS_PER_SM)
    marlin_w4a16_kernel(const Params p, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmA,
                        const __grid_constant__ CUtensorMap tmS, const __grid_constant__ CUtensorMap tmZ) {
  using C = Cfg<scalar_t, NB, NT>;
  constexpr bool F16 = C::F16;
  constexpr int MP = C::MP, STAGES = C::STAGES, NWN = C::NWN, KG = C::KG;
  extern __shared__ uint8_t smem_raw[];

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const bool grouped = p.group_size < p.K;
  const int gps = grouped ? KT / p.group_size : 0;  // groups per 128 k (1, 2, 4)
  // stream-K: the (slice, k stage) list is cut into gridDim.x equal contiguous ranges, so every CTA streams the
  // same number of 16 KiB weight stages whatever the shape; a range may end one slice and start the next
  const int item_begin = (int)((int64_t)blockIdx.x * p.total / gridDim.x);
  const int item_end = (int)((int64_t)(blockIdx.x + 1) * p.total / gridDim.x);
  const int nk = item_end - item_begin;  // > 0: the host keeps grid <= total
  // the 128-byte swizzle pattern is a function of the shared-memory address: align the ring to 1 KiB
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  float* red = reinterpret_cast<float*>(smem + STAGES * C::STAGE_BYTES);
  const uint32_t full_bar = smem_base + STAGES * C::STAGE_BYTES + C::RED_BYTES;
  const uint32_t empty_bar = full_bar + STAGES * 8;
  // fix-up flags of shared slice j (0, 1): [4j] slice, [4j+1] first contributing CTA, [4j+2] contributors, [4j+3] am I the last arriver
  int* s_flag = reinterpret_cast<int*>(smem + STAGES * C::STAGE_BYTES + C::RED_BYTES + 2 * STAGES * 8);

  APH_MMA_STAMP(0, tid == 0);  // entry
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar + s * 8, 1);
      mbar_init(empty_bar + s * 8, CONSUMERS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // programmatic dependent launch: the next kernel of the stream may start filling free CTA slots now; it streams
  // its (immutable) weights while this grid drains and touches activations / outputs only after griddep_wait()
  griddep_launch_dependents();

  if (w == SG * CONSUMERS) {
    // ============================ producer: one thread, tiled TMA loads ============================
    if (lane != 0) {
      if (p.cluster_split > 1) cluster_sync_twice_unaligned();  // every thread takes part in the cluster merge's two phases
      return;
    }
    prefetch_tensormap(&tmB);
    prefetch_tensormap(&tmA);
    if (grouped) prefetch_tensormap(&tmS);
    if (HAS_ZP && grouped) prefetch_tensormap(&tmZ);
    const uint64_t pol_stream = policy_evict_first();  // weights are read once
    const uint64_t pol_keep = policy_evict_last();     // activations are shared by every CTA
    const uint32_t stage_tx =
        (uint32_t)(C::B_BYTES + C::A_BYTES + KG * gps * NT * 2 + (HAS_ZP ? KG * gps * (NT / 8) * 4 : 0));
    // arm the stage's barrier with its whole byte count and fetch what does not depend on earlier kernels
    auto load_weights = [&](int st, int slice, int ks) {
      const uint32_t sB = smem_base + st * C::STAGE_BYTES;
      const uint32_t fb = full_bar + st * 8;
      if (p.dbg == 2) {
        mbar_arrive(fb);
        return;
      }
      mbar_expect_tx(fb, stage_tx);
      tma_load_2d_hint(sB, &tmB, slice * NT, ks * 8 * KG, fb, pol_stream);
      if (grouped) {
        tma_load_2d(sB + C::B_BYTES + C::A_BYTES, &tmS, slice * NT, ks * KG * gps, fb);
        if constexpr (HAS_ZP) tma_load_2d(sB + C::B_BYTES + C::A_BYTES + C::S_BYTES, &tmZ, slice * (NT / 8), ks * KG * gps, fb);
      }
    };
    auto load_activations = [&](int st, int ks) {
      if (p.dbg == 2) return;
      const uint32_t sA = smem_base + st * C::STAGE_BYTES + C::B_BYTES;
#pragma unroll
      for (int i = 0; i < 2 * KG; ++i)
        tma_load_2d_hint(sA + i * C::A_BOX, &tmA, ks * C::KS + 64 * i, p.m0, full_bar + st * 8, pol_keep);
    };
    // the first STAGES items: weights at once (the ring starts empty; item i < STAGES sits in slot i under either ring
    // mapping), their activations after the dependency wait
    int pre_ks[STAGES];
    int it = 0;
    bool waited = false;
    RingCursor<STAGES> ring;  // the producer handles every item
    auto end_prologue = [&]() {
      griddep_wait();  // activations come from the previous kernel
      const int n = it < STAGES ? it : STAGES;
      for (int i = 0; i < n; ++i) load_activations(i, pre_ks[i]);
      waited = true;
    };
#ifdef APH_RING_FAULT
    int held_st = -1, held_slice = 0, held_ks = 0;  // an odd item held back while the even item behind it goes out first
#endif
    for_each_segment(item_begin, item_end, p.k_stages, [&](int slice, int k0, int cnt, bool) {
#pragma unroll 1
      for (int j = 0; j < cnt; ++j, ++it) {
        if (it < STAGES) {
          load_weights(it, slice, k0 + j);
          pre_ks[it] = k0 + j;
          ring.advance(it);
        } else {
          if (!waited) end_prologue();
          const int st = ring.slot(it);
          mbar_wait(empty_bar + st * 8, ring.phase(it) ^ 1);
          ring.advance(it);
#ifdef APH_RING_FAULT
          if (it & 1) {
            held_st = st;
            held_slice = slice;
            held_ks = k0 + j;
            continue;
          }
#endif
          load_weights(st, slice, k0 + j);
          load_activations(st, k0 + j);
#ifdef APH_RING_FAULT
          if (held_st >= 0) {
            ring_fault_delay();
            load_weights(held_st, held_slice, held_ks);
            load_activations(held_st, held_ks);
            held_st = -1;
          }
#endif
        }
      }
    });
    if (!waited) end_prologue();
    APH_MMA_STAMP(1, true);  // producer: last load issued
#ifdef APH_RING_FAULT
    if (held_st >= 0) {
      load_weights(held_st, held_slice, held_ks);
      load_activations(held_st, held_ks);
    }
#endif
    if (p.cluster_split > 1) cluster_sync_twice_unaligned();
    return;
  }

  // ================================ consumer warps ================================
  const int sg = w / CONSUMERS, wi = w % CONSUMERS;
  const int wn = wi % NWN, wg = wi / NWN;
  const int g = lane >> 2, t = lane & 3;
  const uint32_t offB = (uint32_t)(wg * 8 * (NT * 8) + wn * 512 + lane * 16);
  // ldmatrix row address inside a 128B-swizzled activation box: row rr, 16-byte chunk c -> rr*128 + ((c ^ (rr&7)) << 4)
  const uint32_t a_row = (NB == 1) ? (uint32_t)(lane & 7) : (uint32_t)(((lane >> 4) & 1) * 8 + (lane & 7));
  const uint32_t a_x = (uint32_t)(((lane >> 3) & 1) ^ (lane & 7));  // (k-half bit) ^ (row & 7)
  const uint32_t offA = (uint32_t)(C::B_BYTES + wg * 2 * C::A_BOX) + a_row * 128;
  uint32_t a_sw[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) a_sw[i] = ((uint32_t)(i << 1) ^ a_x) << 4;

  // part = sum over the current quantisation group of a_k (q_kn - z_n), exact 16-bit weights times activations in fp32;
  // acc += s_n * part at the end of the group
  float acc[4][NB][4], part[4][NB][4];
  auto clear_accumulators = [&]() {
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int b = 0; b < NB; ++b)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[j][b][e] = part[j][b][e] = 0.f;
  };
  // what the spliced value OFF + q has subtracted from it, as a packed pair per (tile j, row g / g + 8): OFF + 8 for
  // the symmetric types (uint4b8), OFF + zero point otherwise. (OFF + q) - (OFF + z) is exact in f16 and in bf16.
  uint32_t czp[4][2];
#pragma unroll
  for (int j = 0; j < 4; ++j) czp[j][0] = czp[j][1] = C::MAGIC | 0x00080008u;
  auto decode_zp = [&](uint32_t zword) {
    // nibble e' of the packed word holds column block interleave[e'] = {0,2,4,6,1,3,5,7}[e']
    constexpr int inv[8] = {0, 4, 1, 5, 2, 6, 3, 7};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t z0 = (zword >> (4 * inv[2 * j])) & 0xf, z1 = (zword >> (4 * inv[2 * j + 1])) & 0xf;
      czp[j][0] = C::MAGIC | z0 | (z0 << 16);
      czp[j][1] = C::MAGIC | z1 | (z1 << 16);
    }
  };
  // acc += s * part; part = 0.   sc[j][h]: tile j, row g (h = 0) / g + 8 (h = 1)
  auto flush = [&](const float (&sc)[4][2]) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        acc[j][b][0] = fmaf(sc[j][0], part[j][b][0], acc[j][b][0]);
        acc[j][b][1] = fmaf(sc[j][0], part[j][b][1], acc[j][b][1]);
        acc[j][b][2] = fmaf(sc[j][1], part[j][b][2], acc[j][b][2]);
        acc[j][b][3] = fmaf(sc[j][1], part[j][b][3], acc[j][b][3]);
        part[j][b][0] = part[j][b][1] = part[j][b][2] = part[j][b][3] = 0.f;
      }
  };
  auto flush_from_smem = [&](const uint8_t* stage, int gl) {
    const uint8_t* sSp = stage + C::B_BYTES + C::A_BYTES;
    const uint4 sv = *reinterpret_cast<const uint4*>(sSp + gl * (NT * 2) + wn * 128 + g * 16);
    const uint32_t sw[4] = {sv.x, sv.y, sv.z, sv.w};
    float sc[4][2];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if constexpr (F16) {
        const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&sw[j]));
        sc[j][0] = f.x;
        sc[j][1] = f.y;
      } else {
        sc[j][0] = __uint_as_float(sw[j] << 16);
        sc[j][1] = __uint_as_float(sw[j] & 0xffff0000u);
      }
    }
    flush(sc);
  };
  auto sub2 = [](uint32_t x, uint32_t c) -> uint32_t {
    if constexpr (F16) {
      __half2 v = __hsub2(*reinterpret_cast<__half2*>(&x), *reinterpret_cast<__half2*>(&c));
      return *reinterpret_cast<uint32_t*>(&v);
    } else {
      __nv_bfloat162 v = __hsub2(*reinterpret_cast<__nv_bfloat162*>(&x), *reinterpret_cast<__nv_bfloat162*>(&c));
      return *reinterpret_cast<uint32_t*>(&v);
    }
  };

  // k16-rows per quantisation group (2, 4 or 8: powers of two) as a mask / shift; channelwise has one group per segment
  const int rpg_mask = grouped ? p.group_size / 16 - 1 : 0x7fffffff;
  const int rpg_shift = grouped ? 31 - __clz(p.group_size / 16) : 31;

  constexpr int CT = SG * CONSUMERS * 32;
  static_assert(SG == 2, "RingCursor splits the ring between two consumer groups");
  RingCursor<STAGES> ring;  // this warp handles the items of group sg only (it & 1 == sg), in order across segments
  griddep_wait();  // free here (no activation tile arrives before the producer passed its own wait); orders the global writes below

  // ---------------- segments: the pieces of output slices inside this CTA's item range ----------------
  int it0 = 0;        // CTA-local index of the segment's first item (ring position)
  int n_shared = 0;   // shared slices met so far (at most two)
  for_each_segment(item_begin, item_end, p.k_stages, [&](int slice, int k0, int cnt, bool whole) {
    clear_accumulators();
    if constexpr (HAS_ZP) {
      if (!grouped) decode_zp(*reinterpret_cast<const uint32_t*>(p.zp + ((int64_t)slice * (NT / 8) + wn * 8 + g) * 4));
    }

#pragma unroll 1
    for (int it = it0 + ((sg - it0) & (SG - 1)); it < it0 + cnt; it += SG) {
      const int st = ring.slot(it);
      mbar_wait(full_bar + st * 8, ring.phase(it));
      APH_MMA_STAMP(2, tid == 0 && it == 0);  // first stage landed
      ring.advance(it);
      const uint32_t sbase = smem_base + st * C::STAGE_BYTES;
      if (p.dbg != 1)
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        if constexpr (HAS_ZP) {
          // first k16-row of a quantisation group: its zero points
          if ((r & 1) == 0 && (r & rpg_mask) == 0 && grouped)
            decode_zp(*reinterpret_cast<const uint32_t*>(smem + st * C::STAGE_BYTES + C::B_BYTES + C::A_BYTES + C::S_BYTES +
                                                         (wg * gps + (r >> rpg_shift)) * (NT / 8) * 4 + wn * 32 + g * 4));
        }
        uint32_t q[4];
        lds128(sbase + offB + r * (NT * 8), q[0], q[1], q[2], q[3]);
        uint32_t bf[NB][2];
        const uint32_t aaddr = sbase + offA + (r >> 2) * C::A_BOX + a_sw[r & 3];
        if constexpr (NB == 1) {
          ldsm_x2(aaddr, bf[0][0], bf[0][1]);
        } else {
#pragma unroll
          for (int bb = 0; bb < NB; bb += 2) ldsm_x4(aaddr + bb * 8 * 128, bf[bb][0], bf[bb][1], bf[bb + 1][0], bf[bb + 1][1]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          // exponent splice (OFF + q, one LOP3 per pair) and exact subtraction of OFF + z: registers a0 / a2 hold
          // row g of the tile (k low / high half), a1 / a3 row g + 8
          const uint32_t a0 = sub2(and_or(q[j], 0x000f000fu, C::MAGIC), czp[j][0]);
          const uint32_t a2 = sub2(and_or(q[j] >> 4, 0x000f000fu, C::MAGIC), czp[j][0]);
          const uint32_t a1 = sub2(and_or(q[j] >> 8, 0x000f000fu, C::MAGIC), czp[j][1]);
          const uint32_t a3 = sub2(and_or(q[j] >> 12, 0x000f000fu, C::MAGIC), czp[j][1]);
#pragma unroll
          for (int b = 0; b < NB; ++b) mma16816<F16>(part[j][b], a0, a1, a2, a3, bf[b][0], bf[b][1]);
        }
        // quantisation-group boundary: group sizes 32 / 64 end inside the warp block, 128 at its end
        if ((r & 1) == 1 && ((r + 1) & rpg_mask) == 0)
          flush_from_smem(smem + st * C::STAGE_BYTES, wg * gps + (r >> rpg_shift));
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar + st * 8);
    }
    if (!grouped) {
      // channelwise: one scale row, "single" permutation over 32-column blocks (marlin_utils.py:204-213)
      float sc[4][2];
      const scalar_t* srow = reinterpret_cast<const scalar_t*>(p.s) + (int64_t)slice * NT + wn * 64;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int pos = (j >> 1) * 32 + 8 * (g >> 1) + (g & 1) + 4 * (j & 1);
        sc[j][0] = ScalarTraits<scalar_t>::to_float(srow[pos]);
        sc[j][1] = ScalarTraits<scalar_t>::to_float(srow[pos + 2]);
      }
      flush(sc);
    }

    APH_MMA_STAMP(3, tid == 0);  // main loop of the (last) segment done
    // ---------------- reduce the consumer warps through shared memory ----------------
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int b = 0; b < NB; ++b) {
        float* r0 = red + (size_t)((sg * KG + wg) * MP + b * 8 + 2 * t) * C::RP + wn * 64 + 16 * j + g;
        r0[0] = acc[j][b][0];
        r0[C::RP] = acc[j][b][1];
        r0[8] = acc[j][b][2];
        r0[C::RP + 8] = acc[j][b][3];
      }
    named_bar_sync(1, CT);

    const int nelem = p.m_rows * NT;
    if (p.cluster_split > 1) {
      // ---- cluster split-K: the CTAs of this cluster hold the partial sums of ONE slice (aligned plan: a CTA's range is
      //      one piece of one slice). Ranks 1.. push their partial into rank 0's idle pipeline ring through DSMEM, rank 0
      //      adds them in rank order and writes the result: no scratch round trip, no atomics, no last-arriver pass.
      //      Phase one of the cluster barrier = "my main loop is over, my ring is idle" (and: every CTA runs). ----
      const int crank = (int)cluster_ctarank();
      cluster_arrive_release();
      cluster_wait_acquire();
      if (crank != 0) {
        for (int idx = tid; idx < nelem; idx += CT) {
          const int m = idx / NT, n = idx % NT;
          float v = 0.f;
#pragma unroll
          for (int kg = 0; kg < SG * KG; ++kg) v += red[(size_t)(kg * MP + m) * C::RP + n];
          st_cluster_f32(cluster_map_shared(smem_base + (uint32_t)(((crank - 1) * MP * NT + idx) * 4), 0u), v);
        }
        cluster_arrive_release();  // rank 0 reads after this phase
        cluster_wait_acquire();
        it0 += cnt;
        return;
      }
      cluster_arrive_release();
      cluster_wait_acquire();
      const float* recv = reinterpret_cast<const float*>(smem);
      scalar_t* cout = reinterpret_cast<scalar_t*>(p.c) + (int64_t)p.m0 * p.N + (int64_t)slice * NT;
      for (int idx = tid; idx < nelem; idx += CT) {
        const int m = idx / NT, n = idx % NT;
        float v = 0.f;
#pragma unroll
        for (int kg = 0; kg < SG * KG; ++kg) v += red[(size_t)(kg * MP + m) * C::RP + n];
        for (int r = 1; r < p.cluster_split; ++r) v += recv[(r - 1) * MP * NT + idx];  // rank order: deterministic
        cout[(int64_t)m * p.N + n] = ScalarTraits<scalar_t>::from_float(v);
      }
      it0 += cnt;
      return;
    }
    if (whole) {
      // the whole k range of the slice was ours: write the result
      scalar_t* cout = reinterpret_cast<scalar_t*>(p.c) + (int64_t)p.m0 * p.N + (int64_t)slice * NT;
      for (int idx = tid; idx < nelem; idx += CT) {
        const int m = idx / NT, n = idx % NT;
        float v = 0.f;
#pragma unroll
        for (int kg = 0; kg < SG * KG; ++kg) v += red[(size_t)(kg * MP + m) * C::RP + n];
        cout[(int64_t)m * p.N + n] = ScalarTraits<scalar_t>::from_float(v);
      }
      named_bar_sync(1, CT);  // `red` is reused by the next segment
    } else {
      // the slice is shared with neighbouring CTAs: fp32 partial to the scratch, arrival published by one thread
      // (barrier + its fence make the other threads' stores visible first); the sum is deferred to the end
      const int s_begin = slice * p.k_stages;
      float* mine = p.scratch + ((size_t)blockIdx.x * 2 + (item_begin < s_begin ? 1 : 0)) * (size_t)(MP * NT);
      for (int idx = tid; idx < nelem; idx += CT) {
        const int m = idx / NT, n = idx % NT;
        float v = 0.f;
#pragma unroll
        for (int kg = 0; kg < SG * KG; ++kg) v += red[(size_t)(kg * MP + m) * C::RP + n];
        __stcg(mine + idx, v);
      }
      named_bar_sync(1, CT);
      if (tid == 0) {
        // CTAs whose ranges meet [s_begin, s_begin + k_stages): owner(x) = ceil((x + 1) * grid / total) - 1
        const uint32_t G = gridDim.x, T = (uint32_t)p.total;  // (total + 1) * grid < 2^31 (host)
        const int first = (int)((((uint32_t)s_begin + 1u) * G + T - 1u) / T) - 1;
        const int last = (int)((((uint32_t)(s_begin + p.k_stages)) * G + T - 1u) / T) - 1;
        fence_acq_rel_gpu();
        const int arrived = atomicAdd(&p.counters[slice], 1);
        s_flag[4 * n_shared + 0] = slice;
        s_flag[4 * n_shared + 1] = first;
        s_flag[4 * n_shared + 2] = last - first + 1;
        s_flag[4 * n_shared + 3] = (arrived == last - first);
      }
      ++n_shared;
    }
    it0 += cnt;
  });

  APH_MMA_STAMP(4, tid == 0);  // segments written / published
  // ---------------- deferred fix-up: the last arriver of a shared slice sums the partials in CTA order ----------------
  if (n_shared == 0) return;
  named_bar_sync(1, CT);  // flags visible, `red` free
  sum_shared_slices<scalar_t, MP, NT, CT>(reinterpret_cast<scalar_t*>(p.c) + (int64_t)p.m0 * p.N, p.scratch, p.counters, p.N, p.m_rows,
                                          p.k_stages, p.total, s_flag, n_shared, reinterpret_cast<int*>(red), tid);
  APH_MMA_STAMP(5, tid == 0);  // fix-up done
}

}  // namespace mg
}  // namespace aphb200
