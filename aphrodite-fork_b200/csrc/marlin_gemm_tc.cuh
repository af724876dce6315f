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

constexpr int NTILE = 128;   // output columns per CTA (= MMA M)
constexpr int KSTAGE = 128;  // k per pipeline stage (= 8 k16-rows of the Marlin layout)
constexpr int DQ_WARPS = 8;  // two per TMEM lane quadrant
constexpr int THREADS = (3 + DQ_WARPS) * 32;
constexpr int W_BYTES = 8 * NTILE * 8;
constexpr int S_BYTES = 4 * NTILE * 2;
constexpr int Z_BYTES = 4 * (NTILE / 8) * 4;
constexpr int W_STAGE = W_BYTES + S_BYTES + Z_BYTES;  // 9472 bytes

struct Params {
  const uint8_t* s;   // scales (channelwise path reads them directly)
  const uint8_t* zp;  // zero points or null
  void* c;
  float* scratch;     // [n_slices][split][MB*128] fp32 partials (split > 1 only)
  int* counters;      // [n_slices] arrival counters, all-zero on entry and on exit
  int N, K;
  int m0, m_rows;     // this launch covers rows [m0, m0 + m_rows) of a / c
  int group_size;     // == K for channelwise
  int k_stages;       // ceil(K / 128)
  int split;          // CTAs sharing one output slice
  int stages_per_chunk;
};

template <typename scalar_t, int MB>
struct Cfg {
  static constexpr bool F16 = std::is_same<scalar_t, __half>::value;
  static constexpr int ACT_BOX = MB * 128;  // MB rows x 64 k, 128B-swizzled
  static constexpr int ACT_BYTES = 2 * ACT_BOX;
  static constexpr int CTAS_PER_SM = MB <= 64 ? 2 : 1;
  static constexpr int TMEM_COLS = CTAS_PER_SM == 2 ? 256 : 512;
  static constexpr int D_COLS = MB <= 64 ? 64 : 256;
  static constexpr int A_COL0 = D_COLS;
  // activation ring == TMEM weight-slot ring
  static constexpr int SA_TMEM = (TMEM_COLS - D_COLS) / 64;
  static constexpr int BUDGET = (227 * 1024) / CTAS_PER_SM - 1024 - 512 - (CTAS_PER_SM == 2 ? 1024 : 0);
  static constexpr int SA_SMEM = (BUDGET - 3 * W_STAGE) / ACT_BYTES;
  static constexpr int SA = SA_SMEM < SA_TMEM ? SA_SMEM : SA_TMEM;
  static_assert(SA >= 2, "activation ring too shallow");
  static constexpr int SW_FIT = (BUDGET - SA * ACT_BYTES) / W_STAGE;
  static constexpr int SW = SW_FIT > 8 ? 8 : SW_FIT;
  static_assert(SW >= 3, "weight ring too shallow");
  static constexpr int ACT_OFF = 0;                  // activation slots first: 1 KiB aligned
  static constexpr int W_OFF = SA * ACT_BYTES;
  static constexpr int BAR_OFF = W_OFF + SW * W_STAGE;
  static constexpr int BAR_BYTES = 512;
  static constexpr int SMEM_BYTES = 1024 + BAR_OFF + BAR_BYTES;
  static constexpr uint32_t MAGIC = F16 ? 0x64006400u : 0x43004300u;  // 1024 / 128 as packed pairs
  static constexpr uint32_t IDESC = umma_idesc_f16(F16 ? 0 : 1, NTILE, MB);
};

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

template <typename scalar_t, int MB, bool HAS_ZP>
__global__ void __launch_bounds__(THREADS, Cfg<scalar_t, MB>::CTAS_PER_SM)
    marlin_w4a16_tc_kernel(const Params p, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmA,
                           const __grid_constant__ CUtensorMap tmS, const __grid_constant__ CUtensorMap tmZ) {
  using C = Cfg<scalar_t, MB>;
  constexpr bool F16 = C::F16;
  constexpr int SA = C::SA, SW = C::SW;
  extern __shared__ uint8_t smem_raw[];

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const bool grouped = p.group_size < p.K;
  const int gps = grouped ? KSTAGE / p.group_size : 0;  // groups per stage (1, 2, 4)
  const int slice = blockIdx.x / p.split, chunk = blockIdx.x % p.split;
  const int kb = chunk * p.stages_per_chunk;
  const int nk = min(p.k_stages, kb + p.stages_per_chunk) - kb;  // > 0 by construction
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t bar0 = smem_base + C::BAR_OFF;
  const uint32_t fullw_bar = bar0, emptyw_bar = bar0 + 64, fulla_bar = bar0 + 128, ready_bar = bar0 + 192, emptya_bar = bar0 + 256,
                 done_bar = bar0 + 320;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + C::BAR_OFF + 328);
  int* s_last = reinterpret_cast<int*>(smem + C::BAR_OFF + 332);

  if (tid == 0) {
    for (int s = 0; s < SW; ++s) {
      mbar_init(fullw_bar + s * 8, 1);
      mbar_init(emptyw_bar + s * 8, DQ_WARPS);
    }
    for (int s = 0; s < SA; ++s) {
      mbar_init(fulla_bar + s * 8, 1);
      mbar_init(ready_bar + s * 8, DQ_WARPS);
      mbar_init(emptya_bar + s * 8, 1);
    }
    mbar_init(done_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (w == 1) tmem_alloc(smem_u32(tmem_slot), C::TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (w == 0) {
    // ==================== weight producer: packed weights, scales, zero points ====================
    if (lane != 0) return;
    prefetch_tensormap(&tmB);
    if (grouped) prefetch_tensormap(&tmS);
    if (HAS_ZP && grouped) prefetch_tensormap(&tmZ);
    const uint64_t pol_stream = policy_evict_first();  // weights are read once
    const uint32_t stage_tx = (uint32_t)(W_BYTES + gps * NTILE * 2 + (HAS_ZP ? gps * (NTILE / 8) * 4 : 0));
#pragma unroll 1
    for (int it = 0; it < nk; ++it) {
      const int sw = it % SW, ks = kb + it;
      mbar_wait_relaxed(emptyw_bar + sw * 8, ((it / SW) & 1) ^ 1);
      const uint32_t dst = smem_base + C::W_OFF + sw * W_STAGE;
      const uint32_t fb = fullw_bar + sw * 8;
      mbar_expect_tx(fb, stage_tx);
      tma_load_2d_hint(dst, &tmB, slice * NTILE, ks * 8, fb, pol_stream);
      if (grouped) {
        tma_load_2d(dst + W_BYTES, &tmS, slice * NTILE, ks * gps, fb);
        if constexpr (HAS_ZP) tma_load_2d(dst + W_BYTES + S_BYTES, &tmZ, slice * (NTILE / 8), ks * gps, fb);
      }
    }
    return;
  }

  if (w == 2 + DQ_WARPS) {
    // ==================== activation producer ====================
    if (lane != 0) return;
    prefetch_tensormap(&tmA);
    const uint64_t pol_keep = policy_evict_last();  // activations are shared by every CTA
#pragma unroll 1
    for (int it = 0; it < nk; ++it) {
      const int sa = it % SA, ks = kb + it;
      mbar_wait_relaxed(emptya_bar + sa * 8, ((it / SA) & 1) ^ 1);
      const uint32_t dst = smem_base + C::ACT_OFF + sa * C::ACT_BYTES;
      const uint32_t fb = fulla_bar + sa * 8;
      mbar_expect_tx(fb, (uint32_t)C::ACT_BYTES);
      tma_load_2d_hint(dst, &tmA, ks * KSTAGE, p.m0, fb, pol_keep);
      tma_load_2d_hint(dst + C::ACT_BOX, &tmA, ks * KSTAGE + 64, p.m0, fb, pol_keep);
    }
    return;
  }

  if (w == 1) {
    // ============================ MMA issuer: one thread ============================
    if (lane == 0) {
#pragma unroll 1
      for (int it = 0; it < nk; ++it) {
        const int sa = it % SA;
        const uint32_t ph = (it / SA) & 1;
        mbar_wait_relaxed(fulla_bar + sa * 8, ph);  // activations have landed (async proxy writes visible)
        mbar_wait_relaxed(ready_bar + sa * 8, ph);  // dequantised weights are in TMEM
        tc_fence_after();
        const uint32_t act = smem_base + C::ACT_OFF + sa * C::ACT_BYTES;
        const uint32_t a_tmem = tmem_base + C::A_COL0 + sa * 64;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint64_t bdesc = umma_desc_sw128(act + (kk >> 2) * C::ACT_BOX + (kk & 3) * 32);
          tc_mma_ts(tmem_base, a_tmem + kk * 8, bdesc, C::IDESC, (it > 0 || kk > 0) ? 1u : 0u);
        }
        tc_commit(emptya_bar + sa * 8);  // slot (activations + TMEM weights) free once these MMAs retire
      }
      tc_commit(done_bar);
    }
    __syncwarp();
    named_bar_sync(2, (DQ_WARPS + 1) * 32);  // every TMEM read of the epilogue is done
    tmem_dealloc(tmem_base, C::TMEM_COLS);
    return;
  }

  // ================================ dequantiser warps ================================
  const int qd = w & 3;            // TMEM lane quadrant this warp may access
  const int h = (w - 2) >> 2;      // which half of the stage's k16-rows
  const int g = lane >> 2;
  const uint32_t offW = (uint32_t)(C::W_OFF + (4 * h) * (NTILE * 8) + (qd >> 1) * 512 + lane * 16 + (qd & 1) * 8);
  const uint32_t offS = (uint32_t)(C::W_OFF + W_BYTES + (qd >> 1) * 128 + g * 16 + (qd & 1) * 8);
  const uint32_t offZ = (uint32_t)(C::W_OFF + W_BYTES + S_BYTES + (qd >> 1) * 32 + g * 4);
  const uint32_t a_lane = (uint32_t)(32 * qd) << 16;
  const int rpg_shift = grouped ? 31 - __clz(p.group_size / 16) : 3;  // log2(k16-rows per group): 1, 2, 3
  const bool two_groups = grouped && p.group_size == 32;              // a warp's 4 rows span two groups

  uint32_t s_lo[2], s_hi[2], c_lo[2], c_hi[2];  // per 16-row tile: packed (x, x) for rows g and g + 8
  c_lo[0] = c_lo[1] = c_hi[0] = c_hi[1] = C::MAGIC | 0x00080008u;
  s_lo[0] = s_lo[1] = s_hi[0] = s_hi[1] = 0;
  auto set_zp = [&](uint32_t zword) {
    // nibble e' of the packed word holds column block {0,2,4,6,1,3,5,7}[e']: tile j rows g / g+8 sit in
    // nibbles {0,4,1,5,2,6,3,7}[2j] and [2j+1] = j and j + 4
#pragma unroll
    for (int tt = 0; tt < 2; ++tt) {
      const int j = 2 * (qd & 1) + tt;
      const uint32_t zl = (zword >> (4 * j)) & 0xf;
      const uint32_t zh = (zword >> (4 * j + 16)) & 0xf;
      c_lo[tt] = C::MAGIC | zl | (zl << 16);
      c_hi[tt] = C::MAGIC | zh | (zh << 16);
    }
  };
  auto load_group = [&](uint32_t wstage, int gl) {
    uint32_t sw0, sw1;
    lds64(wstage + offS + gl * (NTILE * 2), sw0, sw1);
    s_lo[0] = __byte_perm(sw0, sw0, 0x1010);
    s_hi[0] = __byte_perm(sw0, sw0, 0x3232);
    s_lo[1] = __byte_perm(sw1, sw1, 0x1010);
    s_hi[1] = __byte_perm(sw1, sw1, 0x3232);
    if constexpr (HAS_ZP) set_zp(lds32(wstage + offZ + gl * (NTILE / 8) * 4));
  };
  if (HAS_ZP && !grouped) set_zp(*reinterpret_cast<const uint32_t*>(p.zp + ((int64_t)slice * (NTILE / 8) + (qd >> 1) * 8 + g) * 4));

  int pending = -1;  // activation/TMEM slot whose stores are issued but not yet published to the MMA warp
  auto publish = [&]() {
    tmem_wait_st();
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(ready_bar + pending * 8);
  };

#pragma unroll 1
  for (int it = 0; it < nk; ++it) {
    const int sw = it % SW, sa = it % SA;
    mbar_wait(fullw_bar + sw * 8, (it / SW) & 1);
    const uint32_t wstage = smem_base + sw * W_STAGE;
    uint32_t q[4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i) lds64(wstage + offW + i * (NTILE * 8), q[i][0], q[i][1]);
    uint32_t a[2][16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (grouped && (i == 0 || (i == 2 && two_groups))) load_group(wstage, (4 * h + i) >> rpg_shift);
#pragma unroll
      for (int tt = 0; tt < 2; ++tt) {
        const uint32_t a0 = and_or(q[i][tt], 0x000f000fu, C::MAGIC);
        const uint32_t a2 = and_or(q[i][tt] >> 4, 0x000f000fu, C::MAGIC);
        const uint32_t a1 = and_or(q[i][tt] >> 8, 0x000f000fu, C::MAGIC);
        const uint32_t a3 = and_or(q[i][tt] >> 12, 0x000f000fu, C::MAGIC);
        a[tt][4 * i + 0] = sub_mul2<F16>(a0, c_lo[tt], s_lo[tt], grouped);
        a[tt][4 * i + 1] = sub_mul2<F16>(a1, c_hi[tt], s_hi[tt], grouped);
        a[tt][4 * i + 2] = sub_mul2<F16>(a2, c_lo[tt], s_lo[tt], grouped);
        a[tt][4 * i + 3] = sub_mul2<F16>(a3, c_hi[tt], s_hi[tt], grouped);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(emptyw_bar + sw * 8);  // this warp's words of the stage are in registers
    if (pending >= 0) publish();                      // the previous stage's stores had the math above to land
    mbar_wait(emptya_bar + sa * 8, ((it / SA) & 1) ^ 1);  // the MMAs that read this TMEM slot have retired
    tc_fence_after();
    const uint32_t a_tmem = tmem_base + a_lane + C::A_COL0 + sa * 64 + (4 * h) * 8;
    tmem_st_16x128b_x8(a_tmem, a[0]);
    tmem_st_16x128b_x8(a_tmem + (16u << 16), a[1]);
    pending = sa;
  }
  if (pending >= 0) publish();

  // ================================ epilogue ================================
  mbar_wait(done_bar, 0);
  tc_fence_after();
  const int n_local = 32 * qd + lane;  // TMEM lane == output column inside the slice
  float s_n = 1.f;
  if (!grouped) {
    // channelwise scales use the "single" permutation over 32-column blocks (marlin_utils.py:204-213)
    const int nn = n_local & 63, j = nn >> 4, rr = nn & 15, gg = rr & 7;
    const int pos = (j >> 1) * 32 + 8 * (gg >> 1) + (gg & 1) + 4 * (j & 1) + ((rr >> 3) << 1);
    s_n = ScalarTraits<scalar_t>::to_float(reinterpret_cast<const scalar_t*>(p.s)[(int64_t)slice * NTILE + (n_local & 64) + pos]);
  }
  scalar_t* cout = reinterpret_cast<scalar_t*>(p.c) + (int64_t)p.m0 * p.N + (int64_t)slice * NTILE + n_local;
  float* base = p.scratch + (size_t)slice * p.split * (size_t)(MB * NTILE);
  float* mine = base + (size_t)chunk * (MB * NTILE);
#pragma unroll 1
  for (int cc = h; cc < MB / 16; cc += 2) {
    if (16 * cc >= p.m_rows) break;
    uint32_t v[16];
    tmem_ld_32x32b_x16(tmem_base + a_lane + 16 * cc, v);
    tmem_wait_ld();
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int m = 16 * cc + j;
      if (m < p.m_rows) {
        const float val = __uint_as_float(v[j]) * s_n;
        if (p.split == 1)
          cout[(int64_t)m * p.N] = ScalarTraits<scalar_t>::from_float(val);
        else
          __stcg(mine + m * NTILE + n_local, val);
      }
    }
  }
  tc_fence_before();
  named_bar_arrive(2, (DQ_WARPS + 1) * 32);
  if (p.split == 1) return;

  // ---------------- split-K: the last CTA to arrive sums the partials in chunk order ----------------
  constexpr int CT = DQ_WARPS * 32;
  const int ctid = tid - 64;
  __threadfence();
  named_bar_sync(1, CT);
  if (ctid == 0) *s_last = (atomicAdd(&p.counters[slice], 1) == p.split - 1);
  named_bar_sync(1, CT);
  if (!*s_last) return;
  __threadfence();
  const int nvec = p.m_rows * (NTILE / 4);
  scalar_t* crow = reinterpret_cast<scalar_t*>(p.c) + (int64_t)p.m0 * p.N + (int64_t)slice * NTILE;
  const float4* base4 = reinterpret_cast<const float4*>(base);
  for (int idx = ctid; idx < nvec; idx += CT) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int ci = 0; ci < p.split; ++ci) {
      const float4 t = __ldcg(base4 + (size_t)ci * (MB * NTILE / 4) + idx);
      v.x += t.x;
      v.y += t.y;
      v.z += t.z;
      v.w += t.w;
    }
    const int m = idx / (NTILE / 4), n = (idx % (NTILE / 4)) * 4;
    scalar_t o[4] = {ScalarTraits<scalar_t>::from_float(v.x), ScalarTraits<scalar_t>::from_float(v.y),
                     ScalarTraits<scalar_t>::from_float(v.z), ScalarTraits<scalar_t>::from_float(v.w)};
    *reinterpret_cast<uint2*>(crow + (int64_t)m * p.N + n) = *reinterpret_cast<const uint2*>(o);
  }
  if (ctid == 0) p.counters[slice] = 0;  // leave the caller's lock buffer all-zero
}

}  // namespace mgtc
}  // namespace aphb200
