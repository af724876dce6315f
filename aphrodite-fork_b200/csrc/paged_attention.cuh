// Paged single-query (decode) attention for sm_100a: GQA-packed split-K kernel + reduce.
//
// Replaces /root/reference/kernels/attention/attention_kernels.cuh:86-495 (main kernel) and
// :563-668 (v2 reduce). Same inputs, same result within the stated tolerance; different design:
//
//   * one CTA per (sequence, kv head [x tile of 8 q heads], partition): all q heads that share
//     a kv head are packed into one MMA operand, so every KV byte is fetched from HBM once
//     (the reference launches one CTA per q head and re-reads each kv head G times);
//   * KV pages are streamed by cp.async.bulk (1-D TMA bulk copy, SASS UBLKCP) straight from
//     the paged cache layout into a per-warp ring of shared-memory stages, completion tracked
//     by mbarriers - no register staging, no CTA-wide barriers in the main loop;
//   * S^T = K Q^T and O^T = V^T P^T run on mma.sync.m16n8k16 with TOKENS as the M dimension
//     and the 8 packed q heads as N, so no tensor-core work is wasted on head padding and the
//     cache layouts ([D/x][bs][x] for K, [D][bs] for V) are exactly the row-major A operands;
//   * FP8 KV is converted in registers (cvt.rn.f16x2.e4m3x2 / e5m2x2, exact) and fed to f16
//     MMAs; k_scale is folded into the logit scale and v_scale into the output;
//   * online softmax in the exp2 domain with warp-shuffle max/sum reductions, one pass over KV
//     (the reference makes two sweeps and round-trips logits through shared memory);
//   * the split-K partition size is chosen from the problem size; up to 8 partitions of one (sequence, kv head) form
//     a thread-block CLUSTER and are merged on chip: every CTA pushes its fp32 partial rows into the shared memory of
//     the cluster's CTA that owns the head (DSMEM stores), one cluster barrier, and the owners write `out` - no partial
//     round trip through HBM and no second kernel. Only splits finer than 8 go through fp32 partials in a workspace
//     (one per cluster) and the reduce kernel.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

namespace aphb200 {
namespace pa {

constexpr int NW = 4;            // warps per CTA
constexpr int THREADS = NW * 32;
constexpr int HEADS_PER_CTA = 8;  // q heads packed into the MMA N dimension
constexpr float LOG2E = 1.4426950408889634f;
// Two builds of the main kernel. The default gives every warp a 16 KiB ring of pipeline stages and room for the
// cluster merge of any cluster size up to 8 (rows a CTA receives: one per source rank x owned head, CP * ceil(8 / CP),
// at most 14): three CTAs per SM. The DENSE build trades ring depth for occupancy - 12 KiB rings, 8 receive rows
// (clusters of 1, 2, 4, 8) - so that four CTAs fit an SM; it is ahead where many waves of CTAs flow through the SMs
// (the kernel is then short of warps to hide latencies, not of bytes in flight per warp) and behind on one-wave grids
// (profiles/r02_attention_dense_variant.txt). Used for the 1-byte caches at head sizes 64 and 128.
constexpr int ring_bytes_for(bool dense) { return dense ? 12288 : 16384; }
constexpr int recv_slots_for(int head_size, bool dense = false) { return head_size <= 128 && !dense ? 14 : 8; }

struct Params {
  void* out;                 // [S, H, D] scalar
  const void* q;             // [S, H, D] scalar
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes;  // [H] or null
  const float* k_scale;       // device scalars (fp8 only)
  const float* v_scale;
  float* part_o;              // [S, H, P, D] fp32 partial numerators
  float* part_ml;             // [S, H, P, 2] (max in log2 domain, sum)
  int num_seqs, num_heads, num_kv_heads;
  int max_blocks_per_seq;
  int partition_tokens;       // tokens per partition (multiple of the stage size)
  int num_partitions;         // grid z: partitions per (sequence, kv head tile), a multiple of cluster_parts
  int cluster_parts;          // CTAs per thread-block cluster along z (1..8): the partitions a cluster merges on chip
  int dense;                  // run the four-CTAs-per-SM build of the kernel where there is one (see ring_bytes_for)
  int64_t q_stride, out_stride, kv_block_stride, kv_head_stride;  // elements
  float scale;
  int tp_rank, bs_local_blocks, bs_vert_stride, bs_block_size, bs_head_sliding_step;
};

using namespace ::aphb200::ptx;

// two fp8 (low 16 bits of w, or high 16 bits) -> f16x2, exact.
template <int KV>
__device__ __forceinline__ uint32_t fp8x2_to_f16x2(uint16_t v) {
  uint32_t d;
  if constexpr (KV == APH_KV_FP8_E5M2)
    asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(d) : "h"(v));
  else
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(d) : "h"(v));
  return d;
}

template <bool F16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (F16) {
    __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  } else {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
  }
}

// Token permutation used with FP8 KV (see kernel comment): S^T row r <-> token TAU(r) of the group.
__device__ __forceinline__ int tau_fp8(int r) {
  const int rr = r & 7, tt = rr >> 1;
  const int hi = (r >> 3);                      // 0: rows 0-7, 1: rows 8-15
  const int shift = ((tt >= 2) ^ hi) ? 2 : 0;   // rows<8: +2 for tt>=2 ; rows>=8: +2 for tt<2
  return 4 * tt + shift + (rr & 1);
}

}  // namespace pa
}  // namespace aphb200
