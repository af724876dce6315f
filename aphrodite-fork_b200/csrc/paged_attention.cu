// Host entry points aph_paged_attention_v1 / _v2: argument checks (the reference's TORCH_CHECKs,
// paged_attention_v1.cu:100-168), split-K partition choice and dispatch to the per-head-size
// translation units (paged_attention_inst.cu).
#include <cstdlib>
#include <mutex>

#include "paged_attention.cuh"

// Scratch of aph_paged_attention_v1_noalloc: the split-K partials live in this file-scope device array instead of a
// caller buffer. Every GPU that loads the library holds these bytes until the process exits. In HOST code the name is
// only a stand-in for the symbol, never a device address: the address is looked up per device with
// cudaGetSymbolAddress (static_scratch() below) and that pointer is what the kernels receive.
#define APH_PA_NOALLOC_SCRATCH_BYTES APH_PA_NOALLOC_MAX_SCRATCH_BYTES  // include/aphrodite_b200.h: 16 MiB
static __device__ __align__(256) unsigned char g_pa_noalloc_scratch[APH_PA_NOALLOC_SCRATCH_BYTES];

namespace aphb200 {
namespace pa {

#define APH_DECL(D) int launch_pa_d##D(const Params& p, int dtype, int kv_dtype, int block_size, cudaStream_t stream);
APH_DECL(64)
APH_DECL(128)
#ifndef APH_FAST_BUILD
APH_DECL(32)
APH_DECL(80)
APH_DECL(96)
APH_DECL(112)
APH_DECL(120)
APH_DECL(192)
APH_DECL(256)
#endif
#undef APH_DECL

struct Split {
  int partition_tokens;
  int num_partitions;  // grid z
  int cluster_parts;   // partitions merged on chip by one thread-block cluster (1 = no cluster)
  bool dense;          // the four-CTAs-per-SM build of the kernel (paged_attention.cuh: ring_bytes_for)
  int out_partials() const { return num_partitions / cluster_parts; }  // partials per (sequence, head) in the workspace
};

// Split-K plan. A CTA streams one partition of one (sequence, kv head tile); 3 CTAs share an SM (`slots` = one wave).
// Measured on B200 (tools/bench_attention.py, tools/sweep_attention_split.py; profiles/r02_attention_policy.jsonl,
// r01_attention_split_sweep.log), in "token-times" (the time a CTA of a full wave needs per token, ~15 ns):
//   * a grid that fits ONE wave costs about  partition_tokens + 600  (every CTA pays its start-up and merge once, nothing
//     overlaps them), plus the merge: ~100 when the partitions of a (sequence, kv head) form one thread-block cluster
//     and meet in shared memory, ~250 + 12 per cluster for clusters of 8 followed by the reduce kernel, ~500 for the reduce kernel alone;
//   * a multi-wave grid costs about  (CTAs / slots, rounded up below 3 waves) x (partition_tokens + 370): start-up and tail of one CTA hide
//     under the streaming of its neighbours, so short partitions are cheap - and clusters do NOT pay there (their CTAs
//     are scheduled together and leave together: C2 107 us against 97 us), the partials go through the workspace.
// The cheapest candidate wins. One-wave partitions are multiples of 64 tokens (4 warps x 16-token stages; a multiple
// of every page size). A cluster of CP CTAs needs CP * ceil(heads per CTA / CP) rows in the kernel's receive buffer
// (14 for head sizes up to 128, 8 above). APH_PA_CLUSTER (development): 0 = never use clusters, 1 = whenever the split
// allows, 2 = as described.
static Split choose_split(int num_seqs, int num_heads, int num_kv_heads, int head_size, int max_seq_len, bool has_dense) {
  const int G = num_heads / num_kv_heads;
  const int tiles = (G + HEADS_PER_CTA - 1) / HEADS_PER_CTA;
  const int64_t units = (int64_t)num_seqs * num_kv_heads * tiles;
  const int64_t slots = (int64_t)sm_count() * 3;
  const int msl = max_seq_len < 1 ? 1 : max_seq_len;
  // development overrides (tools/sweep_attention_split.py): waves of CTAs aimed at, smallest partition, cluster policy
  static const char* const e_w0 = getenv("APH_PA_WAVES");  // read once per process unless APH_DEV_ENV=1
  static const char* const e_m0 = getenv("APH_PA_MIN_PART");
  static const char* const e_c0 = getenv("APH_PA_CLUSTER");
  const char* const e_w = dev_env_live() ? getenv("APH_PA_WAVES") : e_w0;
  const char* const e_m = dev_env_live() ? getenv("APH_PA_MIN_PART") : e_m0;
  const char* const e_c = dev_env_live() ? getenv("APH_PA_CLUSTER") : e_c0;
  const int waves = e_w ? atoi(e_w) : 4;
  const int min_part = e_m ? atoi(e_m) : 256;
  const int policy = e_c ? atoi(e_c) : 2;
  const int gt = G < HEADS_PER_CTA ? G : HEADS_PER_CTA;
  const int recv_slots = recv_slots_for(head_size);  // Cfg::RECV_SLOTS
  auto fits = [&](int cp) { return cp * ((gt + cp - 1) / cp) <= recv_slots; };
  static const char* const e_d0 = getenv("APH_PA_DENSE");  // development: 0 never, 1 wherever the build exists
  const char* const e_d = dev_env_live() ? getenv("APH_PA_DENSE") : e_d0;

  Split best{msl, 1, 1, false};
  double best_cost = 1e30;
  auto consider = [&](int pt, int np, int cp, double cost, bool dense = false) {
    if (cost < best_cost) {
      best_cost = cost;
      best = Split{pt, np, cp, dense};
    }
  };
  // ---- multi-wave candidate: aim at `waves` waves, every partition through the workspace ----
  {
    const int64_t target = slots * (waves > 0 ? waves : 4);
    int64_t p_want = units > 0 ? (target + units - 1) / units : 1;
    if (p_want < 1) p_want = 1;
    int64_t pt = (msl + p_want - 1) / p_want;
    pt = ((pt + 255) / 256) * 256;
    if (pt < min_part) pt = (min_part + 255) / 256 * 256;
    const int np = (int)((msl + pt - 1) / pt);
    const double w = (double)units * np / (double)slots;
    int cp = 1, z = np;
    if (policy == 1 && np > 1) {  // development: clusters wherever the split allows
      cp = np <= 8 ? np : 8;
      while (!fits(cp)) ++cp;
      z = (np + cp - 1) / cp * cp;
    }
    // few waves run in rounds (1.15 waves cost two), many waves flow
    const double rounds = w <= 3.0 ? (double)(int64_t)(w + 0.999) : w;
    // the reduce kernel hides under the last wave of a multi-wave grid (~150); behind a single wave it is exposed (~500)
    // many waves: the dense build (4 CTAs per SM, 8 receive rows) is 1.5-2 % ahead (C2 96.7 -> 95.3 us, 256 x 1024 91.7 -> 89.7)
    bool dense = has_dense && w > 3.0 && (cp == 1 || cp * ((gt + cp - 1) / cp) <= recv_slots_for(head_size, true));
    if (e_d) dense = has_dense && atoi(e_d) != 0 && (cp == 1 || cp * ((gt + cp - 1) / cp) <= recv_slots_for(head_size, true));
    consider((int)pt, z, cp, rounds * (double)(pt + (w <= 1.0 ? 600 : 370)) + (np > 1 ? (w <= 1.0 ? 500.0 : 150.0) : 0.0), dense);
  }
  // ---- one-wave candidates ----
  if (e_w == nullptr && e_m == nullptr && units > 0 && units <= slots) {
    const int max_np = (int)(slots / units);
    for (int want = 1; want <= max_np && want <= 512; ++want) {
      int64_t pt = (msl + want - 1) / want;
      pt = ((pt + 63) / 64) * 64;
      const int np = (int)((msl + pt - 1) / pt);
      if ((int64_t)np * units > slots) continue;
      if (np == 1) {
        consider((int)pt, 1, 1, (double)pt + 600.0);
        continue;
      }
      int cp = 1, z = np;
      double merge = 500.0;
      if (policy != 0) {
        if (np <= 8) {
          int c = np;
          while (!fits(c)) ++c;
          if ((int64_t)c * units <= slots) {
            cp = c;
            z = c;  // partitions that start beyond a sequence stay empty
            merge = 100.0;
          }
        } else if ((int64_t)((np + 7) / 8 * 8) * units <= slots) {
          cp = 8;
          z = (np + 7) / 8 * 8;
          merge = 250.0 + 12.0 * (z / 8);  // the reduce kernel walks one partial per cluster (C5 x 1: 16 partials 12.8 us, 32 partials 16.8 us)
        }
      }
      consider((int)pt, z, cp, (double)pt + 600.0 + merge);
    }
  }
  return best;
}

static int run(void* out, const void* query, const void* key_cache, const void* value_cache, int num_seqs,
               int num_heads, int num_kv_heads, int head_size, float scale, const int32_t* block_tables,
               const int32_t* seq_lens, int block_size, int max_seq_len, int max_num_blocks_per_seq,
               const float* alibi_slopes, int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
               int64_t kv_head_stride, int dtype, int kv_dtype, const float* k_scale, const float* v_scale,
               int tp_rank, int bs_local_blocks, int bs_vert_stride, int bs_block_size, int bs_head_sliding_step,
               void* workspace, size_t workspace_bytes, void* stream) {
  APH_CHECK(num_seqs >= 0 && num_heads > 0 && num_kv_heads > 0 && head_size > 0, APH_ERR_INVALID_ARGUMENT,
            "paged_attention: bad sizes");
  APH_CHECK(num_heads % num_kv_heads == 0, APH_ERR_INVALID_ARGUMENT,
            "paged_attention: num_heads %d not a multiple of num_kv_heads %d", num_heads, num_kv_heads);
  APH_CHECK(kv_dtype == APH_KV_AUTO || (k_scale && v_scale), APH_ERR_INVALID_ARGUMENT,
            "paged_attention: fp8 kv cache needs k_scale and v_scale");
  const int esz = kv_dtype == APH_KV_AUTO ? (dtype == APH_DT_F32 ? 4 : 2) : 1;
  APH_CHECK(aligned16(key_cache) && aligned16(value_cache) && (kv_block_stride * esz) % 16 == 0 &&
                (kv_head_stride * esz) % 16 == 0,
            APH_ERR_INVALID_ARGUMENT, "paged_attention: kv cache pages must be 16-byte aligned");
  APH_CHECK(head_size % (16 / esz) == 0, APH_ERR_INVALID_ARGUMENT,
            "paged_attention: head size %d is not a multiple of the %d elements of a key-cache chunk", head_size, 16 / esz);
  APH_CHECK(kv_head_stride == (int64_t)head_size * block_size, APH_ERR_INVALID_ARGUMENT,
            "paged_attention: kv cache must be contiguous within a (block, head) page");
  if (bs_vert_stride > 1)
    APH_CHECK(bs_block_size > 0, APH_ERR_INVALID_ARGUMENT, "paged_attention: blocksparse_block_size must be > 0");
  if (num_seqs == 0) return APH_OK;

  // where paged_attention_inst.cu instantiates the dense build (HAS_DENSE there): 1-byte caches, head sizes 64 and 128,
  // pages of 8 or 16 tokens (a 32-token page is a stage of its own: the ring cannot get shallower)
  const bool has_dense = kv_dtype != APH_KV_AUTO && (head_size == 64 || head_size == 128) && block_size <= 16;
  const Split sp = choose_split(num_seqs, num_heads, num_kv_heads, head_size, max_seq_len, has_dense);
  Params p;
  p.out = out;
  p.q = query;
  p.k_cache = key_cache;
  p.v_cache = value_cache;
  p.block_tables = block_tables;
  p.seq_lens = seq_lens;
  p.alibi_slopes = alibi_slopes;
  p.k_scale = k_scale;
  p.v_scale = v_scale;
  p.num_seqs = num_seqs;
  p.num_heads = num_heads;
  p.num_kv_heads = num_kv_heads;
  p.max_blocks_per_seq = max_num_blocks_per_seq;
  p.partition_tokens = sp.partition_tokens;
  p.num_partitions = sp.num_partitions;
  p.cluster_parts = sp.cluster_parts;
  p.dense = sp.dense ? 1 : 0;
  p.q_stride = q_stride;
  p.out_stride = out_stride;
  p.kv_block_stride = kv_block_stride;
  p.kv_head_stride = kv_head_stride;
  p.scale = scale;
  p.tp_rank = tp_rank;
  p.bs_local_blocks = bs_local_blocks;
  p.bs_vert_stride = bs_vert_stride;
  p.bs_block_size = bs_block_size;
  p.bs_head_sliding_step = bs_head_sliding_step;
  p.part_o = nullptr;
  p.part_ml = nullptr;
  if (sp.out_partials() > 1) {
    const size_t n_o = (size_t)num_seqs * num_heads * sp.out_partials() * head_size;
    const size_t need = n_o * 4 + (size_t)num_seqs * num_heads * sp.out_partials() * 8;
    APH_CHECK(workspace != nullptr && workspace_bytes >= need && aligned16(workspace), APH_ERR_WORKSPACE,
              "paged_attention: workspace of %zu bytes needed, %zu given", need, workspace_bytes);
    p.part_o = (float*)workspace;
    p.part_ml = (float*)workspace + n_o;
  }
  cudaStream_t s = (cudaStream_t)stream;
  switch (head_size) {
    case 64: return launch_pa_d64(p, dtype, kv_dtype, block_size, s);
    case 128: return launch_pa_d128(p, dtype, kv_dtype, block_size, s);
#ifndef APH_FAST_BUILD
    case 32: return launch_pa_d32(p, dtype, kv_dtype, block_size, s);
    case 80: return launch_pa_d80(p, dtype, kv_dtype, block_size, s);
    case 96: return launch_pa_d96(p, dtype, kv_dtype, block_size, s);
    case 112: return launch_pa_d112(p, dtype, kv_dtype, block_size, s);
    case 120: return launch_pa_d120(p, dtype, kv_dtype, block_size, s);
    case 192: return launch_pa_d192(p, dtype, kv_dtype, block_size, s);
    case 256: return launch_pa_d256(p, dtype, kv_dtype, block_size, s);
#endif
    default: return set_error(APH_ERR_INVALID_ARGUMENT, "Unsupported head size: %d", head_size);
  }
}

// Device address of g_pa_noalloc_scratch on the CURRENT device, resolved once per device (set-up; no allocation).
static int static_scratch(void** ptr_out) {
  static std::mutex mu;
  static void* ptrs[64] = {nullptr};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  APH_CHECK(e == cudaSuccess && dev >= 0 && dev < 64, APH_ERR_CUDA, "paged_attention_v1_noalloc: cudaGetDevice: %s",
            cudaGetErrorString(e));
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_pa_noalloc_scratch);
    APH_CHECK(e == cudaSuccess && ptr != nullptr, APH_ERR_CUDA,
              "paged_attention_v1_noalloc: cudaGetSymbolAddress(scratch) on device %d: %s", dev, cudaGetErrorString(e));
    ptrs[dev] = ptr;
  }
  *ptr_out = ptrs[dev];
  return APH_OK;
}

}  // namespace pa
}  // namespace aphb200

using namespace aphb200;

extern "C" size_t aph_paged_attention_workspace_bytes(int num_seqs, int num_heads, int num_kv_heads, int head_size,
                                                      int max_seq_len) {
  if (num_seqs <= 0 || num_heads <= 0 || num_kv_heads <= 0 || head_size <= 0) return 0;
  // (which build of the kernel runs does not change the split, so the cache type need not be known here)
  const pa::Split sp = pa::choose_split(num_seqs, num_heads, num_kv_heads, head_size, max_seq_len, false);
  if (sp.out_partials() <= 1) return 0;
  const size_t n = (size_t)num_seqs * num_heads * sp.out_partials();
  return n * head_size * 4 + n * 8;
}

extern "C" int aph_paged_attention_v1(void* out, const void* query, const void* key_cache, const void* value_cache,
                                      int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                                      const int32_t* block_tables, const int32_t* seq_lens, int block_size,
                                      int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes,
                                      int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                                      int64_t kv_head_stride, int dtype, int kv_dtype, const float* k_scale,
                                      const float* v_scale, int tp_rank, int blocksparse_local_blocks,
                                      int blocksparse_vert_stride, int blocksparse_block_size,
                                      int blocksparse_head_sliding_step, void* workspace, size_t workspace_bytes,
                                      void* stream) {
  return pa::run(out, query, key_cache, value_cache, num_seqs, num_heads, num_kv_heads, head_size, scale,
                 block_tables, seq_lens, block_size, max_seq_len, max_num_blocks_per_seq, alibi_slopes, q_stride,
                 out_stride, kv_block_stride, kv_head_stride, dtype, kv_dtype, k_scale, v_scale, tp_rank,
                 blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                 blocksparse_head_sliding_step, workspace, workspace_bytes, stream);
}

extern "C" int aph_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                                      const void* query, const void* key_cache, const void* value_cache,
                                      int num_seqs, int num_heads, int num_kv_heads, int head_size, float scale,
                                      const int32_t* block_tables, const int32_t* seq_lens, int block_size,
                                      int max_seq_len, int max_num_blocks_per_seq, const float* alibi_slopes,
                                      int64_t q_stride, int64_t out_stride, int64_t kv_block_stride,
                                      int64_t kv_head_stride, int dtype, int kv_dtype, const float* k_scale,
                                      const float* v_scale, int tp_rank, int blocksparse_local_blocks,
                                      int blocksparse_vert_stride, int blocksparse_block_size,
                                      int blocksparse_head_sliding_step, void* workspace, size_t workspace_bytes,
                                      void* stream) {
  (void)exp_sums;
  (void)max_logits;
  (void)tmp_out;  // sized for 512-token partitions by the caller; this implementation partitions differently
  return pa::run(out, query, key_cache, value_cache, num_seqs, num_heads, num_kv_heads, head_size, scale,
                 block_tables, seq_lens, block_size, max_seq_len, max_num_blocks_per_seq, alibi_slopes, q_stride,
                 out_stride, kv_block_stride, kv_head_stride, dtype, kv_dtype, k_scale, v_scale, tp_rank,
                 blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                 blocksparse_head_sliding_step, workspace, workspace_bytes, stream);
}

extern "C" int aph_paged_attention_v1_noalloc(void* out, const void* query, const void* key_cache,
                                              const void* value_cache, int num_seqs, int num_heads, int num_kv_heads,
                                              int head_size, float scale, const int32_t* block_tables,
                                              const int32_t* seq_lens, int block_size, int max_seq_len,
                                              int max_num_blocks_per_seq, const float* alibi_slopes, int64_t q_stride,
                                              int64_t out_stride, int64_t kv_block_stride, int64_t kv_head_stride,
                                              int dtype, int kv_dtype, const float* k_scale, const float* v_scale,
                                              int tp_rank, int blocksparse_local_blocks, int blocksparse_vert_stride,
                                              int blocksparse_block_size, int blocksparse_head_sliding_step,
                                              void* stream) {
  const size_t need = aph_paged_attention_workspace_bytes(num_seqs, num_heads, num_kv_heads, head_size, max_seq_len);
  APH_CHECK(need <= (size_t)APH_PA_NOALLOC_SCRATCH_BYTES, APH_ERR_WORKSPACE,
            "paged_attention_v1_noalloc: this problem needs %zu bytes of split-K scratch, the built-in scratch holds %zu "
            "(use aph_paged_attention_v1 with a caller workspace)",
            need, (size_t)APH_PA_NOALLOC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (need > 0) {
    const int st = pa::static_scratch(&scratch);
    if (st != APH_OK) return st;
  }
  // nothing in the scratch has to start at zero: every partial the reduce kernel reads is written by this call
  return pa::run(out, query, key_cache, value_cache, num_seqs, num_heads, num_kv_heads, head_size, scale,
                 block_tables, seq_lens, block_size, max_seq_len, max_num_blocks_per_seq, alibi_slopes, q_stride,
                 out_stride, kv_block_stride, kv_head_stride, dtype, kv_dtype, k_scale, v_scale, tp_rank,
                 blocksparse_local_blocks, blocksparse_vert_stride, blocksparse_block_size,
                 blocksparse_head_sliding_step, scratch, need > 0 ? (size_t)APH_PA_NOALLOC_SCRATCH_BYTES : 0, stream);
}
