// Main split-K decode attention kernel and the partition reduce kernel (see paged_attention.cuh
// for the design summary).
#pragma once
#include <type_traits>

#include "paged_attention.cuh"

namespace aphb200 {
namespace pa {

template <typename scalar_t, int KV, int D, int BS, bool DENSE = false>
struct Cfg {
  static constexpr int ESZ = (KV == APH_KV_AUTO) ? (int)sizeof(scalar_t) : 1;  // cache element bytes
  static constexpr bool FP8 = KV != APH_KV_AUTO;
  // float32 query with an "auto" cache: the cache holds float32 too (x = 4 per chunk). K and V are rounded to f16 in
  // registers on their way to the MMA (11 significant bits; the reference computes this case in fp32 - the difference
  // is ~5e-4 relative, inside its test's 1e-3 absolute bound for attention-sized values; |K|, |V| must fit f16)
  static constexpr bool F32KV = (KV == APH_KV_AUTO) && std::is_same<scalar_t, float>::value;
  static constexpr bool F16MMA = FP8 || F32KV || std::is_same<scalar_t, __half>::value;
  static constexpr int ST = BS > 16 ? BS : 16;  // tokens per pipeline stage
  static constexpr int SG = ST / 16;            // 16-token MMA groups per stage
  static constexpr int SP = ST / BS;            // cache pages per stage
  static constexpr int KPAGE = D * BS * ESZ;    // bytes of one K (or V) page of one kv head
  static constexpr int STAGE = 2 * SP * KPAGE;  // bytes per stage (K pages then V pages)
  static constexpr int RING = ring_bytes_for(DENSE);
  static constexpr int NS = (RING / STAGE) >= 2 ? ((RING / STAGE) > 8 ? 8 : (RING / STAGE)) : 2;
  static constexpr int NWARPS = (STAGE * NS * 4 <= 160 * 1024) ? 4 : (STAGE * NS * 2 <= 200 * 1024 ? 2 : 1);
  static constexpr int NTHREADS = NWARPS * 32;
  static constexpr int DPAD = (D + 15) / 16 * 16;  // head size 120: the MMAs run over 128, the last 8 are zero (Q) / discarded (O)
  static constexpr int DP = DPAD + 4;  // padded row of the cross-warp merge buffer (floats)
  static constexpr int PIPE_BYTES = NWARPS * NS * STAGE;
  static constexpr int MERGE_BYTES = NWARPS * HEADS_PER_CTA * DP * 4;
  // cluster merge: rows a CTA receives from the partitions of its cluster, [slot][D] fp32 + (max, sum) per slot with
  // slot = source rank x heads this CTA owns; cluster size CP needs CP * ceil(8 / CP) slots (at most 14 for CP <= 8,
  // 8 for the powers of two, which is what head sizes above 128 are limited to so that 3 CTAs still fit an SM)
  static constexpr int RECV_SLOTS = recv_slots_for(D, DENSE);
  static constexpr int RECV_OFF = ((PIPE_BYTES > MERGE_BYTES ? PIPE_BYTES : MERGE_BYTES) + NWARPS * NS * 8 + 15) / 16 * 16;
  static constexpr int RECV_BYTES = RECV_SLOTS * (D + 2) * 4;
  static constexpr int SMEM_BYTES = RECV_OFF + RECV_BYTES;
  static constexpr int CTAS_PER_SM = (SMEM_BYTES + 1024) * 4 <= 227 * 1024 && NTHREADS * 4 * 128 <= 65536
                                         ? 4
                                         : ((SMEM_BYTES + 1024) * 3 <= 227 * 1024 ? 3 : ((SMEM_BYTES + 1024) * 2 <= 227 * 1024 ? 2 : 1));
};

template <typename scalar_t, int KV, int D, int BS, bool DENSE = false>
__global__ void __launch_bounds__(Cfg<scalar_t, KV, D, BS, DENSE>::NTHREADS, Cfg<scalar_t, KV, D, BS, DENSE>::CTAS_PER_SM)
    paged_attention_kernel(const Params p) {
  using C = Cfg<scalar_t, KV, D, BS, DENSE>;
  constexpr bool F16 = C::F16MMA;
  constexpr int NWARPS = C::NWARPS, NS = C::NS, ST = C::ST, SG = C::SG, SP = C::SP, KPAGE = C::KPAGE;
  constexpr int KS = C::DPAD / 16;  // k-steps of Q K^T == m-tiles of P V
  static_assert(D % 16 == 0 || !C::FP8, "an FP8 cache stores 16 elements per chunk: the head size must be a multiple of 16");

  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ float m_s[NWARPS][HEADS_PER_CTA], l_s[NWARPS][HEADS_PER_CTA];

  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  pdl_trigger();  // the next kernel (the split-K reduce) may be set up while this grid drains
  pdl_wait();     // q, the KV cache (the new token was just written) and the block tables come from earlier kernels
  const int G = p.num_heads / p.num_kv_heads;
  const int tiles = (G + HEADS_PER_CTA - 1) / HEADS_PER_CTA;
  const int kvh = blockIdx.x / tiles, tile = blockIdx.x % tiles;
  const int seq = blockIdx.y, part = blockIdx.z;
  const int head0 = kvh * G + tile * HEADS_PER_CTA;       // first q head of this CTA
  const int Gv = min(HEADS_PER_CTA, G - tile * HEADS_PER_CTA);  // valid q heads
  const int L = p.seq_lens[seq];
  const int part_start = part * p.partition_tokens;
  const int P = p.num_partitions;

  // Partitions of one (sequence, kv head tile) are merged inside a thread-block cluster of CP CTAs along z: CTA rank r
  // owns heads [r * HPR, (r + 1) * HPR) of the tile, every CTA pushes its partial rows to the owners (DSMEM), one
  // cluster barrier, the owners finish. First barrier phase = "this CTA runs" (nobody may touch the shared memory of
  // a CTA that has not started).
  const int CP = p.cluster_parts;
  if (CP > 1) cluster_arrive_relaxed();
  const int crank = CP > 1 ? (int)cluster_ctarank() : 0;
  const int HPR = (Gv + CP - 1) / CP;
  float* recv_o = reinterpret_cast<float*>(smem + C::RECV_OFF);
  float* recv_ml = recv_o + C::RECV_SLOTS * D;
  float v_scale = 1.f;
  if constexpr (C::FP8) v_scale = *p.v_scale;

  if (part_start >= L) {
    // Nothing to attend to. With a single partition the output is defined as zeros
    // (reference: empty loops leave accs = 0, attention_kernels.cuh:366-371).
    if (CP == 1) {
      if (P == 1 && L <= 0) {
        scalar_t* o = (scalar_t*)p.out + (int64_t)seq * p.out_stride + (int64_t)head0 * D;
        for (int i = threadIdx.x; i < Gv * D; i += C::NTHREADS) o[i] = ScalarTraits<scalar_t>::from_float(0.f);
      }
      return;
    }
    cluster_wait_acquire();  // end of phase one (this CTA pushes nothing, but may own heads)
  } else {
  // ======================= this CTA's partition =======================
  const int part_end = min(part_start + p.partition_tokens, L);
  const int n_stages = (part_end - part_start + ST - 1) / ST;
  const int n_my = w < n_stages ? (n_stages - w + NWARPS - 1) / NWARPS : 0;
  const int num_seq_pages = (L + BS - 1) / BS;

  uint8_t* my_pipe = smem + (size_t)w * NS * C::STAGE;
  const uint32_t pipe_u32 = smem_u32(my_pipe);
  const uint32_t bar_u32 = smem_u32(smem + (C::PIPE_BYTES > C::MERGE_BYTES ? C::PIPE_BYTES : C::MERGE_BYTES)) + w * NS * 8;
  if (lane == 0) {
#pragma unroll
    for (int s = 0; s < NS; ++s) mbar_init(bar_u32 + s * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // ---- block-table prefetch: lane j holds the page numbers of this warp's stage (ib + j) ----
  const int32_t* bt = p.block_tables + (int64_t)seq * p.max_blocks_per_seq;
  int bt_reg[SP];
  int ib = 0;
  auto bt_fill = [&]() {
    const int si = ib + lane;
#pragma unroll
    for (int pg = 0; pg < SP; ++pg) {
      const int page = (part_start + (w + si * NWARPS) * ST) / BS + pg;
      bt_reg[pg] = (si < n_my && page < num_seq_pages) ? bt[page] : 0;
    }
  };
  bt_fill();

  const uint8_t* kc = (const uint8_t*)p.k_cache + (int64_t)kvh * p.kv_head_stride * C::ESZ;
  const uint8_t* vc = (const uint8_t*)p.v_cache + (int64_t)kvh * p.kv_head_stride * C::ESZ;
  const int64_t blk_bytes = p.kv_block_stride * C::ESZ;

  auto issue = [&](int i) {  // all lanes call (shuffles); lane 0 issues the copies
    if (i - ib >= 32) {
      ib += 32;
      bt_fill();
    }
    int blk[SP];
#pragma unroll
    for (int pg = 0; pg < SP; ++pg) blk[pg] = __shfl_sync(0xffffffffu, bt_reg[pg], i - ib);
    if (lane == 0) {
      const int slot = i % NS;
      const int page0 = (part_start + (w + i * NWARPS) * ST) / BS;
      const int npages = min(SP, num_seq_pages - page0);
      const uint32_t bar = bar_u32 + slot * 8;
      const uint32_t dst = pipe_u32 + slot * C::STAGE;
      mbar_expect_tx(bar, (uint32_t)(npages * 2 * KPAGE));
#pragma unroll
      for (int pg = 0; pg < SP; ++pg) {
        if (pg < npages) {
          const int64_t off = (int64_t)blk[pg] * blk_bytes;
          bulk_g2s(dst + pg * KPAGE, kc + off, KPAGE, bar);
          bulk_g2s(dst + (SP + pg) * KPAGE, vc + off, KPAGE, bar);
        }
      }
    }
  };

  // ---- prologue: the first stages are in flight while the query is fetched and converted ----
#pragma unroll 1
  for (int i = 0; i < NS && i < n_my; ++i) issue(i);

  // ---- Q^T B-fragments (heads along N = lane/4) ----
  float sc = p.scale * LOG2E;
  if constexpr (C::FP8) sc *= *p.k_scale;
  uint32_t qb[KS][2];
  {
    const bool hv = g < Gv;
    const scalar_t* qp = (const scalar_t*)p.q + (int64_t)seq * p.q_stride + (int64_t)(head0 + (hv ? g : 0)) * D;
    float qf[KS][4];
    float amax = 0.f;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int d0 = C::FP8 ? 16 * ks + 4 * t : 16 * ks + 2 * t;
      const int d1 = C::FP8 ? d0 + 2 : d0 + 8;
      const bool v0 = hv && (D % 16 == 0 || d0 < D), v1 = hv && (D % 16 == 0 || d1 < D);
      qf[ks][0] = v0 ? ScalarTraits<scalar_t>::to_float(qp[d0]) : 0.f;
      qf[ks][1] = v0 ? ScalarTraits<scalar_t>::to_float(qp[d0 + 1]) : 0.f;
      qf[ks][2] = v1 ? ScalarTraits<scalar_t>::to_float(qp[d1]) : 0.f;
      qf[ks][3] = v1 ? ScalarTraits<scalar_t>::to_float(qp[d1 + 1]) : 0.f;
      amax = fmaxf(amax, fmaxf(fmaxf(fabsf(qf[ks][0]), fabsf(qf[ks][1])), fmaxf(fabsf(qf[ks][2]), fabsf(qf[ks][3]))));
    }
    float qs = 1.f;
    if constexpr (F16 && !std::is_same<scalar_t, __half>::value) {
      // wide-range query fed to an f16 MMA: pre-scale by a power of two so that it cannot overflow
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
      if (amax > 16384.f && amax < 3.0e38f) {
        int ex;
        frexpf(amax, &ex);
        qs = ldexpf(1.f, 14 - ex);
      }
      sc /= qs;
    }
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      qb[ks][0] = pack2<F16>(qf[ks][0] * qs, qf[ks][1] * qs);
      qb[ks][1] = pack2<F16>(qf[ks][2] * qs, qf[ks][3] * qs);
    }
  }

  // per-thread softmax columns = heads (2t, 2t+1) of this CTA
  float slope[2] = {0.f, 0.f};
  int bs_off[2] = {0, 0};
  const bool sparse = p.bs_vert_stride > 1;
  const int q_bs_id = sparse ? (L - 1) / p.bs_block_size : 0;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const int hh = head0 + min(2 * t + j, Gv - 1);
    if (p.alibi_slopes) slope[j] = p.alibi_slopes[hh] * LOG2E;
    if (sparse)
      bs_off[j] = p.bs_head_sliding_step >= 0 ? (p.tp_rank * p.num_heads + hh) * p.bs_head_sliding_step + 1
                                              : (p.tp_rank * p.num_kv_heads + kvh) * (-p.bs_head_sliding_step) + 1;
  }

  // S^T row r of a 16-token group <-> token rowtok(r) of the group
  const int tokA = C::FP8 ? tau_fp8(g) : g;
  const int tokB = C::FP8 ? tau_fp8(g + 8) : g + 8;

  float o_acc[KS][4];
#pragma unroll
  for (int mt = 0; mt < KS; ++mt)
#pragma unroll
    for (int e = 0; e < 4; ++e) o_acc[mt][e] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

#pragma unroll 1
  for (int i = 0; i < n_my; ++i) {
    const int slot = i % NS;
    mbar_wait(bar_u32 + slot * 8, (i / NS) & 1);
    const uint32_t stage_k = pipe_u32 + slot * C::STAGE;
    const uint32_t stage_v = stage_k + SP * KPAGE;
    const int tok0 = part_start + (w + i * NWARPS) * ST;

#pragma unroll
    for (int sg = 0; sg < SG; ++sg) {
      const int gs = tok0 + sg * 16;
      if (gs >= part_end) break;
      // ---------------- S^T = K Q^T ----------------
      float s[4] = {0.f, 0.f, 0.f, 0.f};
      if constexpr (C::F32KV) {
        // chunks of 4 floats: element d of token ts sits at chunk d / 4, row ts, float d % 4
        const int tsA = sg * 16 + g, tsB = tsA + 8;
        const uint32_t baseA = stage_k + (tsA / BS) * KPAGE + (t >> 1) * (BS * 16) + (tsA % BS) * 16 + (t & 1) * 8;
        const uint32_t baseB = stage_k + (tsB / BS) * KPAGE + (t >> 1) * (BS * 16) + (tsB % BS) * 16 + (t & 1) * 8;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t f[8];
          lds64(baseA + (4 * ks) * (BS * 16), f[0], f[1]);
          lds64(baseB + (4 * ks) * (BS * 16), f[2], f[3]);
          lds64(baseA + (4 * ks + 2) * (BS * 16), f[4], f[5]);
          lds64(baseB + (4 * ks + 2) * (BS * 16), f[6], f[7]);
          uint32_t a[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) a[e] = pack2<true>(__uint_as_float(f[2 * e]), __uint_as_float(f[2 * e + 1]));
          if (D % 16 != 0 && ks == KS - 1) a[2] = a[3] = 0u;
          mma16816<F16>(s, a[0], a[1], a[2], a[3], qb[ks][0], qb[ks][1]);
        }
      } else if constexpr (!C::FP8) {
        const int mi = lane >> 3, r = lane & 7;
        const int ts = sg * 16 + r + (mi & 1) * 8;
        const uint32_t base = stage_k + (ts / BS) * KPAGE + (mi >> 1) * (BS * 16) + (ts % BS) * 16;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t a0, a1, a2, a3;
          ldsm_x4(base + ks * (BS * 32), a0, a1, a2, a3);
          if (D % 16 != 0 && ks == KS - 1) a2 = a3 = 0u;  // elements 120..127 do not exist: whatever lies there must not meet even a zero
          mma16816<F16>(s, a0, a1, a2, a3, qb[ks][0], qb[ks][1]);
        }
      } else {
        const int tsA = sg * 16 + tokA, tsB = sg * 16 + tokB;
        const uint32_t baseA = stage_k + (tsA / BS) * KPAGE + (tsA % BS) * 16 + 4 * t;
        const uint32_t baseB = stage_k + (tsB / BS) * KPAGE + (tsB % BS) * 16 + 4 * t;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          const uint32_t wa = lds32(baseA + ks * (BS * 16));
          const uint32_t wb = lds32(baseB + ks * (BS * 16));
          const uint32_t a0 = fp8x2_to_f16x2<KV>((uint16_t)(wa & 0xffffu));
          const uint32_t a2 = fp8x2_to_f16x2<KV>((uint16_t)(wa >> 16));
          const uint32_t a1 = fp8x2_to_f16x2<KV>((uint16_t)(wb & 0xffffu));
          const uint32_t a3 = fp8x2_to_f16x2<KV>((uint16_t)(wb >> 16));
          mma16816<F16>(s, a0, a1, a2, a3, qb[ks][0], qb[ks][1]);
        }
      }
      // ---------------- online softmax (exp2 domain) ----------------
      const int ta = gs + tokA, tb = gs + tokB;
      float x[4];
      x[0] = s[0] * sc + slope[0] * (float)(ta - L + 1);
      x[1] = s[1] * sc + slope[1] * (float)(ta - L + 1);
      x[2] = s[2] * sc + slope[0] * (float)(tb - L + 1);
      x[3] = s[3] * sc + slope[1] * (float)(tb - L + 1);
      if (ta >= L) x[0] = x[1] = -INFINITY;
      if (tb >= L) x[2] = x[3] = -INFINITY;
      if (sparse) {
        const int kba = (ta / BS) * BS / p.bs_block_size, kbb = (tb / BS) * BS / p.bs_block_size;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const bool keep_a = ((kba + bs_off[j]) % p.bs_vert_stride == 0) || (kba > q_bs_id - p.bs_local_blocks);
          const bool keep_b = ((kbb + bs_off[j]) % p.bs_vert_stride == 0) || (kbb > q_bs_id - p.bs_local_blocks);
          if (!keep_a) x[j] = -INFINITY;
          if (!keep_b) x[2 + j] = -INFINITY;
        }
      }
      float mx[2] = {fmaxf(x[0], x[2]), fmaxf(x[1], x[3])};
#pragma unroll
      for (int o = 4; o <= 16; o <<= 1) {
        mx[0] = fmaxf(mx[0], __shfl_xor_sync(0xffffffffu, mx[0], o));
        mx[1] = fmaxf(mx[1], __shfl_xor_sync(0xffffffffu, mx[1], o));
      }
      float alpha[2], pr[4];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float m_new = fmaxf(m_run[j], mx[j]);
        const float m_use = m_new == -INFINITY ? 0.f : m_new;
        alpha[j] = ex2(m_run[j] - m_use);
        pr[j] = ex2(x[j] - m_use);
        pr[2 + j] = ex2(x[2 + j] - m_use);
        l_run[j] = l_run[j] * alpha[j] + pr[j] + pr[2 + j];
        m_run[j] = m_new;
      }
      if (__any_sync(0xffffffffu, alpha[0] != 1.f || alpha[1] != 1.f)) {
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          o_acc[mt][0] *= alpha[0];
          o_acc[mt][1] *= alpha[1];
          o_acc[mt][2] *= alpha[0];
          o_acc[mt][3] *= alpha[1];
        }
      }
      // P^T B-fragments: transpose the two 8x8 (token x head) blocks across the warp
      const uint32_t pb0 = movmatrix_trans(pack2<F16>(pr[0], pr[1]));
      const uint32_t pb1 = movmatrix_trans(pack2<F16>(pr[2], pr[3]));

      // ---------------- O^T += V^T P^T ----------------
      const bool tail = gs + 16 > L;  // V of tokens >= L may hold anything (NaN): zero it
      if constexpr (C::F32KV) {
        // V page [D][bs] floats: row d, tokens (2t, 2t+1) and (2t+8, 2t+9) of the group are two 8-byte loads
        const int ts0 = sg * 16 + 2 * t, ts1 = ts0 + 8;
        const uint32_t base0 = stage_v + (ts0 / BS) * KPAGE + g * (BS * 4) + (ts0 % BS) * 4;
        const uint32_t base1 = stage_v + (ts1 / BS) * KPAGE + g * (BS * 4) + (ts1 % BS) * 4;
        const int c0 = gs + 2 * t, c2 = c0 + 8;
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          uint32_t f[8];
          lds64(base0 + (16 * mt) * (BS * 4), f[0], f[1]);      // a0: row g
          lds64(base0 + (16 * mt + 8) * (BS * 4), f[2], f[3]);  // a1: row g + 8
          lds64(base1 + (16 * mt) * (BS * 4), f[4], f[5]);      // a2: row g, tokens + 8
          lds64(base1 + (16 * mt + 8) * (BS * 4), f[6], f[7]);  // a3
          if (tail) {
            if (c0 >= L) f[0] = f[2] = 0u;
            if (c0 + 1 >= L) f[1] = f[3] = 0u;
            if (c2 >= L) f[4] = f[6] = 0u;
            if (c2 + 1 >= L) f[5] = f[7] = 0u;
          }
          uint32_t a[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) a[e] = pack2<true>(__uint_as_float(f[2 * e]), __uint_as_float(f[2 * e + 1]));
          mma16816<F16>(o_acc[mt], a[0], a[1], a[2], a[3], pb0, pb1);
        }
      } else if constexpr (!C::FP8) {
        const int mi = lane >> 3, r = lane & 7;
        const int ts = sg * 16 + (mi >> 1) * 8;
        const uint32_t base = stage_v + (ts / BS) * KPAGE + (r + (mi & 1) * 8) * (BS * 2) + (ts % BS) * 2;
        uint32_t mk0 = 0xffffffffu, mk2 = 0xffffffffu;
        if (tail) {
          const int c0 = gs + 2 * t, c2 = gs + 8 + 2 * t;
          mk0 = (c0 < L ? 0x0000ffffu : 0u) | (c0 + 1 < L ? 0xffff0000u : 0u);
          mk2 = (c2 < L ? 0x0000ffffu : 0u) | (c2 + 1 < L ? 0xffff0000u : 0u);
        }
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          uint32_t a0, a1, a2, a3;
          ldsm_x4(base + mt * (16 * BS * 2), a0, a1, a2, a3);
          if (tail) {
            a0 &= mk0;
            a1 &= mk0;
            a2 &= mk2;
            a3 &= mk2;
          }
          mma16816<F16>(o_acc[mt], a0, a1, a2, a3, pb0, pb1);
        }
      } else {
        const int ts = sg * 16 + 4 * t;
        const uint32_t base = stage_v + (ts / BS) * KPAGE + g * BS + (ts % BS);
        uint32_t mk = 0xffffffffu;
        if (tail) {
          const int c = gs + 4 * t;
          mk = (c < L ? 0xffu : 0u) | (c + 1 < L ? 0xff00u : 0u) | (c + 2 < L ? 0xff0000u : 0u) |
               (c + 3 < L ? 0xff000000u : 0u);
        }
        const uint32_t sel = t >= 2 ? 0x1032u : 0x3210u;  // swap 16-bit halves for t >= 2 (see tau_fp8)
#pragma unroll
        for (int mt = 0; mt < KS; ++mt) {
          uint32_t wa = lds32(base + mt * (16 * BS));
          uint32_t wb = lds32(base + mt * (16 * BS) + 8 * BS);
          if (tail) {
            wa &= mk;
            wb &= mk;
          }
          wa = __byte_perm(wa, 0, sel);
          wb = __byte_perm(wb, 0, sel);
          const uint32_t a0 = fp8x2_to_f16x2<KV>((uint16_t)(wa & 0xffffu));
          const uint32_t a2 = fp8x2_to_f16x2<KV>((uint16_t)(wa >> 16));
          const uint32_t a1 = fp8x2_to_f16x2<KV>((uint16_t)(wb & 0xffffu));
          const uint32_t a3 = fp8x2_to_f16x2<KV>((uint16_t)(wb >> 16));
          mma16816<F16>(o_acc[mt], a0, a1, a2, a3, pb0, pb1);
        }
      }
    }
    __syncwarp();
    if (i + NS < n_my) issue(i + NS);
  }

  // ---- per-warp row sums, then cross-warp merge through shared memory ----
#pragma unroll
  for (int o = 4; o <= 16; o <<= 1) {
    l_run[0] += __shfl_xor_sync(0xffffffffu, l_run[0], o);
    l_run[1] += __shfl_xor_sync(0xffffffffu, l_run[1], o);
  }
  __syncthreads();  // every warp is done with its pipeline stages: reuse them as the merge buffer
  float* o_s = reinterpret_cast<float*>(smem);
  if (g == 0) {
    m_s[w][2 * t] = m_run[0];
    m_s[w][2 * t + 1] = m_run[1];
    l_s[w][2 * t] = l_run[0];
    l_s[w][2 * t + 1] = l_run[1];
  }
#pragma unroll
  for (int mt = 0; mt < KS; ++mt) {
    float* r0 = o_s + (size_t)(w * HEADS_PER_CTA + 2 * t) * C::DP + 16 * mt + g;
    r0[0] = o_acc[mt][0];
    r0[C::DP] = o_acc[mt][1];
    r0[8] = o_acc[mt][2];
    r0[C::DP + 8] = o_acc[mt][3];
  }
  __syncthreads();

  if (CP > 1) cluster_wait_acquire();  // end of phase one: every CTA of the cluster runs, its shared memory exists
  const uint32_t recv_o_u32 = smem_u32(recv_o), recv_ml_u32 = smem_u32(recv_ml);
  for (int idx = threadIdx.x; idx < Gv * D; idx += C::NTHREADS) {
    const int h = idx / D, d = idx % D;
    float m_tot = -INFINITY;
#pragma unroll
    for (int ww = 0; ww < NWARPS; ++ww) m_tot = fmaxf(m_tot, m_s[ww][h]);
    const float m_use = m_tot == -INFINITY ? 0.f : m_tot;
    float acc = 0.f, l_tot = 0.f;
#pragma unroll
    for (int ww = 0; ww < NWARPS; ++ww) {
      const float wgt = ex2(m_s[ww][h] - m_use);
      acc += o_s[(size_t)(ww * HEADS_PER_CTA + h) * C::DP + d] * wgt;
      l_tot += l_s[ww][h] * wgt;
    }
    const int head = head0 + h;
    if (CP > 1) {
      // push this partition's row of head h to the CTA of the cluster that owns the head
      const uint32_t owner = (uint32_t)(h / HPR);
      const int slot = crank * HPR + h % HPR;
      st_cluster_f32(cluster_map_shared(recv_o_u32 + (uint32_t)(slot * D + d) * 4u, owner), acc);
      if (d == 0) {
        st_cluster_f32(cluster_map_shared(recv_ml_u32 + (uint32_t)slot * 8u, owner), m_tot);
        st_cluster_f32(cluster_map_shared(recv_ml_u32 + (uint32_t)slot * 8u + 4u, owner), l_tot);
      }
    } else if (P == 1) {
      // reference normaliser: 1 / (sum + 1e-6), attention_kernels.cuh:342
      const float inv = __fdividef(1.f, l_tot + 1e-6f);
      ((scalar_t*)p.out)[(int64_t)seq * p.out_stride + (int64_t)head * D + d] =
          ScalarTraits<scalar_t>::from_float(acc * inv * v_scale);
    } else {
      const int64_t pi = ((int64_t)seq * p.num_heads + head) * P + part;
      p.part_o[pi * D + d] = acc;
      if (d == 0) {
        p.part_ml[pi * 2] = m_tot;
        p.part_ml[pi * 2 + 1] = l_tot;
      }
    }
  }
  }  // this CTA's partition
  if (CP == 1) return;

  // ======================= cluster merge =======================
  cluster_arrive_release();                 // phase two: my rows are on their way
  const int nh = min(HPR, Gv - crank * HPR);  // heads this CTA owns
  if (nh <= 0) return;                      // an exited CTA counts as arrived
  cluster_wait_acquire();                   // every partition's rows have landed here
  const int first_part = part - crank;
  const int np_c = L <= first_part * p.partition_tokens
                       ? 0
                       : min(CP, (L - first_part * p.partition_tokens + p.partition_tokens - 1) / p.partition_tokens);
  const int n_out = P / CP;                 // clusters per (sequence, kv head tile): > 1 = second level through the workspace
  for (int idx = threadIdx.x; idx < nh * D; idx += C::NTHREADS) {
    const int hl = idx / D, d = idx % D;
    float m_tot = -INFINITY;
    for (int src = 0; src < np_c; ++src) m_tot = fmaxf(m_tot, recv_ml[(src * HPR + hl) * 2]);
    const float m_use = m_tot == -INFINITY ? 0.f : m_tot;
    float acc = 0.f, l_tot = 0.f;
    for (int src = 0; src < np_c; ++src) {
      const int slot = src * HPR + hl;
      const float wgt = ex2(recv_ml[slot * 2] - m_use);
      acc += recv_o[slot * D + d] * wgt;
      l_tot += recv_ml[slot * 2 + 1] * wgt;
    }
    const int head = head0 + crank * HPR + hl;
    if (n_out == 1) {
      const float inv = __fdividef(1.f, l_tot + 1e-6f);
      ((scalar_t*)p.out)[(int64_t)seq * p.out_stride + (int64_t)head * D + d] =
          ScalarTraits<scalar_t>::from_float(acc * inv * v_scale);
    } else {
      const int64_t pi = ((int64_t)seq * p.num_heads + head) * n_out + part / CP;
      p.part_o[pi * D + d] = acc;
      if (d == 0) {
        p.part_ml[pi * 2] = m_tot;
        p.part_ml[pi * 2 + 1] = l_tot;
      }
    }
  }
}

// Split-K reduce: out[s,h,:] = sum_j O_j 2^(m_j - m) / (sum_j l_j 2^(m_j - m) + 1e-6) * v_scale.
// Follows the reference's reduce (attention_kernels.cuh:604-667) on fp32 partials.
template <typename scalar_t, int D>
__global__ void __launch_bounds__(128) paged_attention_reduce_kernel(const Params p, const bool fp8) {
  // launched with programmatic stream serialization right behind the main kernel: everything it reads is that
  // kernel's output, so wait for it first (the launch latency is what overlaps)
  pdl_wait();
  pdl_trigger();
  constexpr int WCAP = 512;  // partition weights kept in shared memory (recomputed beyond)
  __shared__ float s_w[WCAP];
  __shared__ float s_red[2][4];
  const int head = blockIdx.x, seq = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, nw = blockDim.x >> 5;
  const int L = p.seq_lens[seq];
  const int P = p.num_partitions;
  const int np = L <= 0 ? 0 : min(P, (L + p.partition_tokens - 1) / p.partition_tokens);
  const int64_t base = ((int64_t)seq * p.num_heads + head) * P;
  const float2* ml = reinterpret_cast<const float2*>(p.part_ml) + base;  // (max, sum) of every partition
  const float v_scale = fp8 ? *p.v_scale : 1.f;
  // the threads split the partitions: global max, then weights 2^(m_j - m) and the denominator
  float m_loc = -INFINITY;
  for (int j = tid; j < np; j += blockDim.x) m_loc = fmaxf(m_loc, ml[j].x);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m_loc = fmaxf(m_loc, __shfl_xor_sync(0xffffffffu, m_loc, o));
  if (lane == 0) s_red[0][w] = m_loc;
  __syncthreads();
  float m_tot = s_red[0][0];
  for (int i = 1; i < nw; ++i) m_tot = fmaxf(m_tot, s_red[0][i]);
  const float m_use = m_tot == -INFINITY ? 0.f : m_tot;
  float d_loc = 0.f;
  for (int j = tid; j < np; j += blockDim.x) {
    const float2 v = ml[j];
    const float wj = ex2(v.x - m_use);
    if (j < WCAP) s_w[j] = wj;
    d_loc += v.y * wj;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d_loc += __shfl_xor_sync(0xffffffffu, d_loc, o);
  if (lane == 0) s_red[1][w] = d_loc;
  __syncthreads();
  float denom = 0.f;
  for (int i = 0; i < nw; ++i) denom += s_red[1][i];
  const float inv = __fdividef(1.f, denom + 1e-6f) * v_scale;
  const float* po = p.part_o + base * D;
  for (int d = tid; d < D; d += blockDim.x) {
    float acc = 0.f;
    int j = 0;
    for (; j + 8 <= np; j += 8) {  // eight independent loads in flight
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = po[(int64_t)(j + u) * D + d];
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += v[u] * (j + u < WCAP ? s_w[j + u] : ex2(ml[j + u].x - m_use));
    }
    for (; j < np; ++j) acc += po[(int64_t)j * D + d] * (j < WCAP ? s_w[j] : ex2(ml[j].x - m_use));
    ((scalar_t*)p.out)[(int64_t)seq * p.out_stride + (int64_t)head * D + d] =
        ScalarTraits<scalar_t>::from_float(acc * inv);
  }
}

}  // namespace pa
}  // namespace aphb200
