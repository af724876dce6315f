// W4A16 dequant-GEMM on Marlin-format weights for sm_100a, small/medium batch path.
//
// Replaces /root/reference/kernels/quantization/gptq_marlin/gptq_marlin.cu (Marlin<> :513-1765,
// marlin_mm :2075-2259, entry :2263-2468) for 4-bit weights. It reads the reference's wire format
// as-is but is organised differently:
//
//   * the weight tile is the A operand of mma.sync.m16n8k16 (16 output columns x 16 k) and the
//     batch is the N=8 dimension: one packed int32 of the Marlin layout IS one A fragment
//     (nibble pairs (0,4) (2,6) (1,5) (3,7) -> a0 a1 a2 a3), so small batches waste no
//     tensor-core rows and the group scales are per accumulator ROW (8 per thread, one 16-byte
//     load straight from the Marlin-permuted scale layout);
//   * int4 -> bf16/f16: exponent splice ((q & 0xf) | 0x4300 = 128 + q, one LOP3 per pair) and one exact packed
//     subtraction of (128 + zero point) - 11 ALU ops per 8 weights (reference: 15). The group scale is NOT applied per
//     weight: the MMA accumulates sum_k a_k (q_kn - z_n) in fp32 over one quantisation group and the scale multiplies
//     that sum once (acc += s[g,n] * part), so no product is rounded to 16 bits. (Round 1 fed the un-subtracted
//     128 + q to the MMA and cancelled the offset on the accumulator; for f16, where the offset is 1024, the tensor
//     core's truncating fp32 accumulation then biased results by ~1 % for activations with a large mean.)
//   * stream-K: the slice-major list of (output slice of 64/128/256 columns, 16 KiB weight stage) items is cut
//     into one equal contiguous range per co-resident CTA slot (2 per SM), so every SM streams the same number
//     of bytes whatever the shape (224 slices on 296 slots left 72 SMs half idle before); a slice shared by
//     neighbouring CTAs meets in an fp32 scratch with a deterministic "last arriver sums in CTA order" fix-up
//     (no spin locks; the caller's lock buffer is the per-slice arrival counter and is left all-zero);
//   * a dedicated producer thread streams weights, activations, scales and zero points with 2-D tiled TMA loads
//     (SASS UTMALDG) into a ring of stages, full/empty mbarriers, no CTA-wide barrier in the main loop; each
//     consumer warp owns 64 output columns and whole 128-row stages (two warp groups take alternate stages), so
//     the per-group flush is amortised over 8 k16 steps;
//   * programmatic dependent launch: the kernel triggers its dependents at once and the next GEMM of the stream
//     fills its ring with weights while this one drains; activations / outputs are touched after griddepcontrol.wait.
// Large batches (M >= 64) are meant for the tcgen05 path; here they run as chunks of 32 rows.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda.h>  // CUtensorMap types only; the encoder is fetched through the runtime (no libcuda link)

#include "marlin_gemm_kernel.cuh"
#include "marlin_gemm_tc.cuh"
#include "marlin_actorder.cuh"

namespace aphb200 {
namespace mg {

struct Plan {
  int nt, grid, k_stages, n_slices, total;
  int cluster_split;  // aligned split-K plans with 2..8 pieces per slice run as thread-block clusters (0 otherwise)
};

// Decomposition. The output is cut into slices of NT columns and K into stages of 16 KiB of packed weights (so the
// number of (slice, stage) items, N*K/32768, does not depend on NT); the kernel gives CTA i the i-th of `grid` equal
// contiguous ranges of the slice-major item list. Two families of plans come out of that one kernel:
//   aligned   grid = n_slices * split with split | k_stages: every range is one chunk of one slice (plain split-K,
//             no fix-up at all when split = 1);
//   stream-K  grid = all co-resident CTA slots: perfectly even bytes per SM, but ranges straddle slices (two
//             fix-ups per CTA, segment turn-arounds).
// make_plan scores every candidate with a small cost model calibrated on B200 (tools/sweep_gemm_plan.py,
// profiles/r01_gemm_plan_sweep.jsonl): a CTA retires one 16 KiB stage per ~0.75 us whether or not it shares its SM
// (it is bound by its own warps' latency, not by the SM), so t = 0.75 us x (stages of the longest CTA) + 2.0 us if
// slices are shared + 1.8 us if ranges are unaligned (0.75 / 1.8 become 1.3 / 6.0 at 9..16 rows, where every
// segment turn-around moves twice the accumulators); the cheapest candidate wins. APH_GEMM_NT / APH_GEMM_GRID force a plan (development).
// development overrides: read once per process (no getenv on the call path), unless the process was started with
// APH_DEV_ENV=1 (the sweep tools change them between calls)
static int env_int(const char* name) {
  const char* e = getenv(name);
  return e ? atoi(e) : 0;
}
static int forced_nt_env() {
  static const int v = env_int("APH_GEMM_NT");
  return dev_env_live() ? env_int("APH_GEMM_NT") : v;
}
static int forced_grid_env() {
  static const int v = env_int("APH_GEMM_GRID");
  return dev_env_live() ? env_int("APH_GEMM_GRID") : v;
}

// Co-resident thread-block clusters of `size` CTAs of the register kernel <NB, NT> (cudaOccupancyMaxActiveClusters,
// cached): a cluster plan must fit one wave, and clusters pack worse than single CTAs (they cannot straddle GPCs).
template <int NB, int NT>
static int mma_max_clusters_of(int size) {
  static int cache[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  if (cache[size] == 0) {
    using C = Cfg<__nv_bfloat16, NB, NT>;
    auto kern = marlin_w4a16_kernel<__nv_bfloat16, NB, NT, false>;
    int n = -1;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES) == cudaSuccess) {
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3((unsigned)(size * 32));
      cfg.blockDim = dim3(THREADS);
      cfg.dynamicSmemBytes = C::SMEM_BYTES;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = (unsigned)size;
      at[0].val.clusterDim.y = 1;
      at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) n = -1;
    }
    (void)cudaGetLastError();
    cache[size] = n;
  }
  return cache[size];
}
template <int NB>
static int mma_max_clusters_nb(int nt, int size) {
  return nt == 256 ? mma_max_clusters_of<NB, 256>(size) : nt == 128 ? mma_max_clusters_of<NB, 128>(size) : mma_max_clusters_of<NB, 64>(size);
}
static int mma_max_clusters(int nb, int nt, int size) {
  return nb == 1 ? mma_max_clusters_nb<1>(nt, size) : nb == 2 ? mma_max_clusters_nb<2>(nt, size) : mma_max_clusters_nb<4>(nt, size);
}

static bool cluster_allowed() {  // APH_GEMM_CLUSTER=0 (development): split-K pieces always meet through the scratch
  static const bool on = [] {
    const char* e = getenv("APH_GEMM_CLUSTER");
    return e == nullptr || e[0] != '0';
  }();
  return dev_env_live() ? !(getenv("APH_GEMM_CLUSTER") && getenv("APH_GEMM_CLUSTER")[0] == '0') : on;
}

static Plan make_plan(int64_t N, int64_t K, int nb) {
  const int slots = sm_count() * (nb <= 2 ? 2 : 1);
  const int forced_nt = forced_nt_env();
  const int forced_grid = forced_grid_env();
  Plan best{};
  double best_cost = 1e30;
  const int nts[3] = {256, 128, 64};
  for (int i = 0; i < 3; ++i) {
    const int nt = nts[i];
    if (N % nt != 0) continue;
    if (forced_nt != 0 && forced_nt != nt && N % forced_nt == 0) continue;
    const int ks = KT * (CONSUMERS / (nt / 64));
    Plan pl{};
    pl.nt = nt;
    pl.k_stages = (int)((K + ks - 1) / ks);
    pl.n_slices = (int)(N / nt);
    pl.total = pl.n_slices * pl.k_stages;
    // wider slices move fewer activation bytes per weight byte (good at M <= 8); at M = 9..16 their larger
    // reduction tile costs pipeline stages instead
    const double width_bias = (nb == 1 ? -0.1 : 0.3) * (nt == 256 ? 2 : nt == 128 ? 1 : 0);
    auto consider = [&](int grid, bool aligned) {
      if (grid < 1 || grid > pl.total) return;
      const int per_cta = (pl.total + grid - 1) / grid;
      const bool shared = !(aligned && grid == pl.n_slices);
      const int split = aligned ? grid / pl.n_slices : 0;
      // the pieces of a slice meet in shared memory when they can form a thread-block cluster (2..8 CTAs): cheaper than
      // the scratch + counter + last-arriver fix-up of the other shared plans
      // (rank 0 receives the other pieces' partials in its idle pipeline ring: at most 40 KiB of it)
      const bool cluster = cluster_allowed() && aligned && split >= 2 && split <= 8 && (split - 1) * 8 * nb * nt * 4 <= 40 * 1024 &&
                           mma_max_clusters(nb, nt, split) >= pl.n_slices;
      const double t_stage = nb == 1 ? 0.75 : nb == 2 ? 1.3 : 2.2;
      const double cost = t_stage * per_cta + (shared ? (cluster ? 0.8 : 2.0) : 0.0) + (aligned ? 0.0 : (nb == 1 ? 1.8 : 6.0)) + width_bias;
      if (cost < best_cost) {
        best_cost = cost;
        best = pl;
        best.grid = grid;
        best.cluster_split = cluster ? split : 0;
      }
    };
    for (int split = 1; split <= 8; ++split)
      if (pl.k_stages % split == 0 && (int64_t)pl.n_slices * split <= slots) consider(pl.n_slices * split, true);
    consider(pl.total < slots ? pl.total : slots, false);
  }
  if (forced_grid > 0) {
    best.grid = forced_grid < best.total ? forced_grid : best.total;
    best.cluster_split = 0;
  }
  return best;
}

static size_t plan_scratch_bytes(const Plan& pl, int nb) { return (size_t)pl.grid * 2 * (8 * nb) * pl.nt * sizeof(float); }

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      sym = nullptr;
    return reinterpret_cast<EncodeTiledFn>(sym);
  }();
  return fn;
}

// 2-D row-major tensor [rows][cols] of `elt_bytes`-wide elements, row pitch in bytes, box [box_rows][box_cols].
// A tensor map is a pure function of these arguments, and a serving process calls with the same few (weight, scale,
// activation buffer) descriptions over and over: encoded maps are kept in a small per-thread direct-mapped cache, so an
// eager (non-graph) call pays for the four driver calls (~1 us each) only the first time it sees a tensor.
struct MapKey {
  const void* base;
  uint64_t cols, rows, pitch;
  uint32_t box_cols, box_rows;
  int dt, swz, promo;
  bool operator==(const MapKey& o) const {
    return base == o.base && cols == o.cols && rows == o.rows && pitch == o.pitch && box_cols == o.box_cols && box_rows == o.box_rows &&
           dt == o.dt && swz == o.swz && promo == o.promo;
  }
};
struct MapCacheEntry {
  MapKey key;
  bool valid;
  alignas(64) CUtensorMap tm;
};
static int make_map_2d(CUtensorMap* tm, CUtensorMapDataType dt, int elt_bytes, const void* base, uint64_t cols, uint64_t rows,
                       uint64_t pitch_bytes, uint32_t box_cols, uint32_t box_rows, CUtensorMapSwizzle swz,
                       CUtensorMapL2promotion promo) {
  constexpr int CACHE = 256;
  static thread_local MapCacheEntry* cache = nullptr;
  if (cache == nullptr) cache = new MapCacheEntry[CACHE]();  // per thread, lives as long as the thread
  const MapKey key{base, cols, rows, pitch_bytes, box_cols, box_rows, (int)dt, (int)swz, (int)promo};
  uint64_t h = (uint64_t)(uintptr_t)base * 0x9E3779B97F4A7C15ull;
  h ^= (cols * 0x100000001B3ull) ^ (rows << 17) ^ ((uint64_t)box_rows << 40) ^ ((uint64_t)box_cols << 48) ^ ((uint64_t)swz << 3) ^ pitch_bytes;
  MapCacheEntry& e = cache[(h ^ (h >> 29)) % CACHE];
  if (e.valid && e.key == key) {
    *tm = e.tm;
    return APH_OK;
  }
  EncodeTiledFn enc = encode_tiled_fn();
  APH_CHECK(enc != nullptr, APH_ERR_CUDA, "gptq_marlin_gemm: driver has no cuTensorMapEncodeTiled");
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {pitch_bytes};
  const cuuint32_t box[2] = {box_cols, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  (void)elt_bytes;
  const CUresult r = enc(tm, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz, promo,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  APH_CHECK(r == CUDA_SUCCESS, APH_ERR_CUDA,
            "gptq_marlin_gemm: cuTensorMapEncodeTiled failed (%d) for [%llu x %llu] pitch %llu box [%u x %u]", (int)r,
            (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)pitch_bytes, box_rows, box_cols);
  e.key = key;
  e.tm = *tm;
  e.valid = true;
  return APH_OK;
}

struct Operands {
  const void *a, *b, *s, *zp;
  int64_t M, lda;
  int64_t num_groups;
  bool f16;
};

template <typename scalar_t, int NB, int NT, bool HAS_ZP>
static int launch(const Params& p, const Operands& op, int grid, cudaStream_t stream) {
  const int cluster_split = p.cluster_split;
  using C = Cfg<scalar_t, NB, NT>;
  auto kern = marlin_w4a16_kernel<scalar_t, NB, NT, HAS_ZP>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES) != cudaSuccess)
      return set_error(APH_ERR_CUDA, "gptq_marlin_gemm: cannot reserve %d bytes of shared memory", C::SMEM_BYTES);
    attr_set[dev] = true;
  }
  const bool grouped = p.group_size < p.K;
  const int gps = grouped ? KT / p.group_size : 0;
  const CUtensorMapDataType dt16 = op.f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  alignas(64) CUtensorMap tmB, tmA, tmS, tmZ;
  int st = make_map_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, op.b, (uint64_t)p.N, (uint64_t)p.K / 16, (uint64_t)p.N * 8, NT,
                       8 * C::KG, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
  if (st != APH_OK) return st;
  st = make_map_2d(&tmA, dt16, 2, op.a, (uint64_t)p.K, (uint64_t)op.M, (uint64_t)op.lda * 2, 64, C::MP, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  if (st != APH_OK) return st;
  if (grouped) {
    st = make_map_2d(&tmS, dt16, 2, op.s, (uint64_t)p.N, (uint64_t)op.num_groups, (uint64_t)p.N * 2, NT, C::KG * gps,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
    if (st != APH_OK) return st;
    if (HAS_ZP) {
      st = make_map_2d(&tmZ, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, op.zp, (uint64_t)p.N / 8, (uint64_t)op.num_groups,
                       (uint64_t)p.N / 2, NT / 8, C::KG * gps, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
      if (st != APH_OK) return st;
    } else {
      tmZ = tmS;
    }
  } else {
    tmS = tmB;  // never dereferenced
    tmZ = tmB;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (cluster_split > 1) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = (unsigned)cluster_split;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl_enabled()) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
#ifdef APH_GEMM_TRACE
  // development build only: APH_GEMM_TRACE=<file> appends a per-CTA timeline summary of this launch (synchronises!)
  static const char* trace_path = getenv("APH_GEMM_TRACE");
  Params pt = p;
  if (trace_path != nullptr) {
    cudaMalloc(&pt.trace, (size_t)grid * 8 * sizeof(unsigned long long));
    cudaMemsetAsync(pt.trace, 0, (size_t)grid * 8 * sizeof(unsigned long long), stream);
  }
  const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, pt, tmB, tmA, tmS, tmZ);
  if (trace_path != nullptr && le == cudaSuccess) {
    std::vector<unsigned long long> h((size_t)grid * 8);
    cudaStreamSynchronize(stream);
    cudaMemcpy(h.data(), pt.trace, h.size() * 8, cudaMemcpyDeviceToHost);
    cudaFree(pt.trace);
    if (FILE* f = fopen(trace_path, "a")) {
      unsigned long long t0 = ~0ull;
      for (int b = 0; b < grid; ++b)
        if (h[b * 8] && h[b * 8] < t0) t0 = h[b * 8];
      fprintf(f, "# mma kernel NB=%d NT=%d grid=%d total=%d k_stages=%d cluster=%d STAGES=%d: ns after the first CTA's entry, min / median / max over CTAs\n",
              NB, NT, grid, p.total, p.k_stages, cluster_split, C::STAGES);
      const char* names[6] = {"entry", "last load issued", "first stage landed", "main loop done", "segments written", "fix-up done"};
      for (int e = 0; e < 6; ++e) {
        std::vector<long long> v;
        for (int b = 0; b < grid; ++b)
          if (h[b * 8 + e]) v.push_back((long long)(h[b * 8 + e] - t0));
        if (v.empty()) continue;
        std::sort(v.begin(), v.end());
        fprintf(f, "%-20s n=%4zu  %7lld %7lld %7lld\n", names[e], v.size(), v.front(), v[v.size() / 2], v.back());
      }
      fclose(f);
    }
  }
#else
  const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, p, tmB, tmA, tmS, tmZ);
#endif
  APH_CHECK(le == cudaSuccess, APH_ERR_CUDA, "gptq_marlin_gemm: launch failed: %s", cudaGetErrorString(le));
  return APH_OK;
}

template <typename scalar_t, int NB>
static int launch_nb(const Params& p, const Operands& op, const Plan& pl, bool has_zp, cudaStream_t stream) {
  switch (pl.nt) {
    case 256:
      return has_zp ? launch<scalar_t, NB, 256, true>(p, op, pl.grid, stream) : launch<scalar_t, NB, 256, false>(p, op, pl.grid, stream);
    case 128:
      return has_zp ? launch<scalar_t, NB, 128, true>(p, op, pl.grid, stream) : launch<scalar_t, NB, 128, false>(p, op, pl.grid, stream);
    default:
      return has_zp ? launch<scalar_t, NB, 64, true>(p, op, pl.grid, stream) : launch<scalar_t, NB, 64, false>(p, op, pl.grid, stream);
  }
}

template <typename scalar_t>
static int run_typed(Params p, const Operands& op, bool has_zp, float* scratch, size_t scratch_bytes, cudaStream_t stream) {
  const int64_t M = op.M;
  for (int64_t m0 = 0; m0 < M; m0 += 32) {
    const int rows = (int)(M - m0 < 32 ? M - m0 : 32);
    const int nb = rows <= 8 ? 1 : (rows <= 16 ? 2 : 4);
    const Plan pl = make_plan(p.N, p.K, nb);
    p.m0 = (int)m0;
    p.m_rows = rows;
    p.k_stages = pl.k_stages;
    p.total = pl.total;
    p.cluster_split = pl.cluster_split;
    const size_t need = plan_scratch_bytes(pl, nb);
    APH_CHECK(((int64_t)pl.total + pl.k_stages + 1) * pl.grid < (1LL << 31), APH_ERR_UNSUPPORTED,
              "gptq_marlin_gemm: %d work items are beyond the 32-bit scheduling arithmetic", pl.total);
    APH_CHECK(scratch != nullptr && scratch_bytes >= need, APH_ERR_WORKSPACE,
              "gptq_marlin_gemm: scratch of %zu bytes needed, %zu given", need, scratch_bytes);
    p.scratch = scratch;
    int st;
    if (nb == 1)
      st = launch_nb<scalar_t, 1>(p, op, pl, has_zp, stream);
    else if (nb == 2)
      st = launch_nb<scalar_t, 2>(p, op, pl, has_zp, stream);
    else
      st = launch_nb<scalar_t, 4>(p, op, pl, has_zp, stream);
    if (st != APH_OK) return st;
  }
  return APH_OK;
}

// ------------------------------------------------------------------------------------------------
// tcgen05 path (marlin_gemm_tc.cuh): slices of 128 columns x k chunks, batch tiles of up to 256 rows
// ------------------------------------------------------------------------------------------------
struct PlanTc {
  int grid, total, k_stages, m_tiles, n_slices;
  int cluster_split;  // 0 = stream-K; S >= 2 = thread-block clusters of S CTAs split the k range of one tile each
};

// Stream-K: one persistent CTA per SM, each taking an equal share of the (tile, k stage) list; fewer CTAs
// only when the problem is too small to give each at least 4 stages.
// How many thread-block clusters of `size` CTAs of a kernel can be resident at once (cudaOccupancyMaxActiveClusters; the
// CTAs of a cluster share a GPC, so this is NOT #SMs / size: 48 clusters of 3 one-SM CTAs do not fit a B200 although
// 144 <= 148, and a second wave doubles the time). Cached per (kernel, size).
using MaxClustersFn = int (*)(int size);

static PlanTc make_plan_tc(int64_t M, int64_t N, int64_t K, int mb, MaxClustersFn max_clusters = nullptr) {
  PlanTc pl;
  pl.n_slices = (int)(N / mgtc::NTILE);
  pl.k_stages = (int)((K + mgtc::KSTAGE - 1) / mgtc::KSTAGE);
  pl.m_tiles = (int)((M + mb - 1) / mb);
  pl.total = pl.n_slices * pl.m_tiles * pl.k_stages;
  int g = sm_count();
  if (g > pl.total / 4) g = pl.total / 4;
  if (g < 1) g = 1;
  pl.grid = g;
  pl.cluster_split = 0;
  // Few tiles (the qkv / o / down shapes of a decode step, every tensor-parallel shard): stream-K would cut every tile
  // into pieces that meet through an fp32 scratch, an atomic counter and a last-arriver merge - about 5 us of the ~15 us
  // such a call takes (stage traces, profiles/r02_gemm_tc_stage_traces.txt). Instead S CTAs of one thread-block cluster
  // split a tile's k range and add their accumulators through distributed shared memory. APH_TC_CLUSTER=0: never.
  static const bool allow = [] {
    const char* e = getenv("APH_TC_CLUSTER");
    return e == nullptr || e[0] != '0';
  }();
  const int tiles = pl.n_slices * pl.m_tiles;
  if (allow && max_clusters != nullptr && tiles * 2 <= sm_count() && pl.k_stages >= 4) {
    int s = sm_count() / tiles;
    if (s > 8) s = 8;                                   // portable cluster size
    while (s > 1 && (pl.k_stages / s < 2 || max_clusters(s) < tiles)) --s;  // >= 2 stages per CTA, one wave of clusters
    if (s >= 2) {
      pl.cluster_split = s;
      pl.grid = tiles * s;
    }
  }
  return pl;
}

static int tile_rows_tc(int64_t rows) { return rows <= 16 ? 16 : rows <= 32 ? 32 : rows <= 64 ? 64 : rows <= 128 ? 128 : 256; }
static int tile_rows_tc8(int64_t rows) { return rows <= 16 ? 16 : rows <= 64 ? 64 : 256; }  // 8-bit weights: fewer instantiations

static size_t scratch_bytes_tc(int64_t M, int64_t N, int64_t K) {
  size_t worst = 0;
  for (int mb : {tile_rows_tc(M), tile_rows_tc8(M)}) {
    const PlanTc pl = make_plan_tc(M, N, K, mb);
    const size_t need = (size_t)pl.grid * 2 * mb * mgtc::NTILE * sizeof(float);
    if (need > worst) worst = need;
  }
  return worst;
}

template <typename scalar_t, int MB, bool HAS_ZP, int BITS = 4, bool ZPF = false, bool G128 = false>
static int launch_tc(mgtc::Params p, const Operands& op, float* scratch, size_t scratch_bytes, cudaStream_t stream) {
  using C = mgtc::Cfg<scalar_t, MB, BITS>;
  constexpr int WPC = BITS / 4;    // 8-byte words per weight column and k16-row
  constexpr int PACK = 32 / BITS;  // zero points per int32
  auto kern = mgtc::marlin_w4a16_tc_kernel<scalar_t, MB, HAS_ZP, BITS, ZPF, G128>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES) != cudaSuccess)
      return set_error(APH_ERR_CUDA, "gptq_marlin_gemm: cannot reserve %d bytes of shared memory", C::SMEM_BYTES);
    attr_set[dev] = true;
  }
  // co-resident clusters of this instantiation, per cluster size (queried once per size)
  static int max_cl[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  struct Q {
    static int get(int size) {
      if (max_cl[size] == 0) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)(size * 16));
        cfg.blockDim = dim3(mgtc::THREADS);
        cfg.dynamicSmemBytes = C::SMEM_BYTES;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)size;
        at[0].val.clusterDim.y = 1;
        at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, mgtc::marlin_w4a16_tc_kernel<scalar_t, MB, HAS_ZP, BITS, ZPF, G128>, &cfg) != cudaSuccess || n <= 0) {
          (void)cudaGetLastError();
          n = -1;
        }
        max_cl[size] = n;
      }
      return max_cl[size];
    }
  };
  const PlanTc pl = make_plan_tc(op.M, p.N, p.K, MB, &Q::get);
  p.k_stages = pl.k_stages;
  p.m_tiles = pl.m_tiles;
  p.total = pl.total;
  p.M = (int)op.M;
  p.cluster_split = pl.cluster_split;
  const size_t need = pl.cluster_split ? 0 : (size_t)pl.grid * 2 * MB * mgtc::NTILE * sizeof(float);
  APH_CHECK(need == 0 || (scratch != nullptr && scratch_bytes >= need), APH_ERR_WORKSPACE,
            "gptq_marlin_gemm: scratch of %zu bytes needed, %zu given", need, scratch_bytes);
  p.scratch = scratch;
  const bool grouped = p.group_size < p.K;
  const int gps = grouped ? mgtc::KSTAGE / p.group_size : 0;
  const CUtensorMapDataType dt16 = op.f16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  alignas(64) CUtensorMap tmB, tmA, tmS, tmZ;
  int st = make_map_2d(&tmB, CU_TENSOR_MAP_DATA_TYPE_UINT64, 8, op.b, (uint64_t)p.N * WPC, (uint64_t)p.K / 16, (uint64_t)p.N * 8 * WPC,
                       mgtc::NTILE * WPC, 8, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
  if (st != APH_OK) return st;
  st = make_map_2d(&tmA, dt16, 2, op.a, (uint64_t)p.K, (uint64_t)op.M, (uint64_t)op.lda * 2, 64, MB, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
  if (st != APH_OK) return st;
  if (grouped) {
    st = make_map_2d(&tmS, dt16, 2, op.s, (uint64_t)p.N, (uint64_t)op.num_groups, (uint64_t)p.N * 2, mgtc::NTILE, gps,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
    if (st != APH_OK) return st;
    if (HAS_ZP && ZPF) {  // float zero points: a second table shaped like the scales
      st = make_map_2d(&tmZ, dt16, 2, op.zp, (uint64_t)p.N, (uint64_t)op.num_groups, (uint64_t)p.N * 2, mgtc::NTILE, gps,
                       CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
      if (st != APH_OK) return st;
    } else if (HAS_ZP) {
      st = make_map_2d(&tmZ, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, op.zp, (uint64_t)p.N / PACK, (uint64_t)op.num_groups,
                       (uint64_t)p.N / PACK * 4, mgtc::NTILE / PACK, gps, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
      if (st != APH_OK) return st;
    } else {
      tmZ = tmS;
    }
  } else {
    tmS = tmB;  // never dereferenced
    tmZ = tmB;
  }
  p.trace = nullptr;
#ifdef APH_GEMM_TRACE
  // development build only (APH_GEMM_TRACE_BUILD=1): APH_GEMM_TRACE=<file> dumps per-stage clock64 stamps of CTA 0.
  // Allocates and synchronises - none of this exists in the product build.
  static const char* trace_path = getenv("APH_GEMM_TRACE");
  if (trace_path != nullptr) {
    cudaMalloc(&p.trace, 72 * 8 * sizeof(unsigned long long));
    cudaMemsetAsync(p.trace, 0, 72 * 8 * sizeof(unsigned long long), stream);
  }
#endif
  {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)pl.grid);
    cfg.blockDim = dim3(mgtc::THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (pl.cluster_split) {
      attr[na].id = cudaLaunchAttributeClusterDimension;
      attr[na].val.clusterDim.x = (unsigned)pl.cluster_split;
      attr[na].val.clusterDim.y = 1;
      attr[na].val.clusterDim.z = 1;
      ++na;
    }
    if (pdl_enabled()) {
      attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    const cudaError_t le = cudaLaunchKernelEx(&cfg, kern, p, tmB, tmA, tmS, tmZ);
    APH_CHECK(le == cudaSuccess, APH_ERR_CUDA, "gptq_marlin_gemm: launch failed: %s", cudaGetErrorString(le));
  }
#ifdef APH_GEMM_TRACE
  if (trace_path != nullptr) {
    unsigned long long host[72 * 8];
    cudaStreamSynchronize(stream);
    cudaMemcpy(host, p.trace, sizeof(host), cudaMemcpyDeviceToHost);
    cudaFree(p.trace);
    if (FILE* f = fopen(trace_path, "w")) {
      fprintf(f, "# MB=%d grid=%d total=%d SA=%d SW=%d: it mma_start fulla ready issued | dq_start fullw math_done slot_free\n", MB, pl.grid,
              pl.total, C::SA, C::SW);
      const unsigned long long t0 = host[4] ? host[4] : host[0];
      for (int i = 0; i < 64 && host[i * 8 + 3]; ++i) {
        fprintf(f, "%2d", i);
        for (int j = 0; j < 8; ++j) fprintf(f, " %8lld", host[i * 8 + j] ? (long long)(host[i * 8 + j] - t0) : -1LL);
        fprintf(f, "\n");
      }
      fprintf(f, "# epilogue per segment: wait_done done drained counted | then: reductions_start reductions_end setup_done kernel_entry\n");
      for (int i = 64; i < 72; ++i) {
        fprintf(f, "e%d", i - 64);
        for (int j = 0; j < 8; ++j) fprintf(f, " %8lld", host[i * 8 + j] ? (long long)(host[i * 8 + j] - t0) : -1LL);
        fprintf(f, "\n");
      }
      fclose(f);
    }
  }
#endif
  return APH_OK;
}

template <typename scalar_t>
static int run_typed_tc(const Params& p0, const Operands& op0, bool has_zp, int bits, float* scratch, size_t scratch_bytes,
                        cudaStream_t stream) {
  // One launch covers every 256-row batch tile (the item list is (slice, batch tile, k stage)); APH_TC_CHUNK=1 runs
  // one launch per 256 rows instead (development).
  static const bool chunked = getenv("APH_TC_CHUNK") != nullptr;
  const int64_t step = chunked ? 256 : (1LL << 40);
  for (int64_t m0 = 0; m0 < op0.M; m0 += step) {
    const int64_t rows = op0.M - m0 < step ? op0.M - m0 : step;
    Operands op = op0;
    op.a = static_cast<const char*>(op0.a) + (size_t)m0 * (size_t)op0.lda * 2;
    op.M = rows;
    mgtc::Params p{};
    p.s = p0.s;
    p.zp = p0.zp;
    p.c = static_cast<char*>(p0.c) + (size_t)m0 * (size_t)p0.N * 2;
    p.counters = p0.counters;  // arrival counters in the caller's lock buffer (N / 64 * 16 ints): one per slice
    p.N = p0.N;
    p.K = p0.K;
    p.group_size = p0.group_size;
    APH_CHECK((rows + 255) / 256 <= 32, APH_ERR_UNSUPPORTED, "gptq_marlin_gemm: size_m above 8192 is not supported in one call");
    int st;
    if (bits == 0) {
      // HQQ: uint4 weights with float16 zero points (is_zp_float; f16 only, grouped only - checked by the caller)
      if constexpr (std::is_same<scalar_t, __half>::value) {
        const int mb = tile_rows_tc8(rows);
        if (mb == 16)
          st = launch_tc<scalar_t, 16, true, 4, true>(p, op, scratch, scratch_bytes, stream);
        else if (mb == 64)
          st = launch_tc<scalar_t, 64, true, 4, true>(p, op, scratch, scratch_bytes, stream);
        else
          st = launch_tc<scalar_t, 256, true, 4, true>(p, op, scratch, scratch_bytes, stream);
      } else {
        st = set_error(APH_ERR_INVALID_ARGUMENT, "Computation type must be float16 (half) when using float zero points.");
      }
      if (st != APH_OK) return st;
      continue;
    }
    if (bits == 8) {
      // 8-bit weights (uint8b128 / uint8): batch tiles of 16, 64 or 256 rows
      const int mb = tile_rows_tc8(rows);
#define APH_TC_CASE8(MBV)                                                                                 \
  st = has_zp ? launch_tc<scalar_t, MBV, true, 8>(p, op, scratch, scratch_bytes, stream)                  \
              : launch_tc<scalar_t, MBV, false, 8>(p, op, scratch, scratch_bytes, stream);
      if (mb == 16) {
        APH_TC_CASE8(16)
      } else if (mb == 64) {
        APH_TC_CASE8(64)
      } else {
        APH_TC_CASE8(256)
      }
#undef APH_TC_CASE8
      if (st != APH_OK) return st;
      continue;
    }
    // group_size 128 = one group per 128-k stage has a build of its own (mgtc: G128)
    const bool g128 = p.group_size == 128 && p.group_size < p.K;
#define APH_TC_CASE(MBV)                                                                                      \
  case MBV:                                                                                                   \
    if (g128)                                                                                                 \
      st = has_zp ? launch_tc<scalar_t, MBV, true, 4, false, true>(p, op, scratch, scratch_bytes, stream)     \
                  : launch_tc<scalar_t, MBV, false, 4, false, true>(p, op, scratch, scratch_bytes, stream);   \
    else                                                                                                      \
      st = has_zp ? launch_tc<scalar_t, MBV, true>(p, op, scratch, scratch_bytes, stream)                     \
                  : launch_tc<scalar_t, MBV, false>(p, op, scratch, scratch_bytes, stream);                   \
    break;
    switch (tile_rows_tc(rows)) {
      APH_TC_CASE(16)
      APH_TC_CASE(32)
      APH_TC_CASE(64)
      APH_TC_CASE(128)
      APH_TC_CASE(256)
      default:
        st = set_error(APH_ERR_INVALID_ARGUMENT, "gptq_marlin_gemm: no batch tile for %lld rows", (long long)rows);
        break;
    }
#undef APH_TC_CASE
    if (st != APH_OK) return st;
  }
  return APH_OK;
}

// which path serves a problem: APH_GEMM_PATH = mma | tc forces one (development), default by batch size
static bool use_tc_path(int64_t M, int64_t N) {
  if (N % mgtc::NTILE != 0) return false;
  static const int forced = [] {
    const char* e = getenv("APH_GEMM_PATH");
    if (e == nullptr) return 0;
    return e[0] == 't' ? 2 : e[0] == 'm' ? 1 : 0;
  }();
  if (forced) return forced == 2;
  // measured crossover (tools/bench_gemm.py, C3 shapes): the register-accumulator kernel wins up to 16 rows
  return M > 16;
}

// ScalarType ids, kernels/core/scalar_type.hpp:100-138 (values pinned in tests/golden/marlin_golden.npz)
constexpr int64_t ID_UINT4 = 1125899906843648LL;
constexpr int64_t ID_UINT4B8 = 1125899907892224LL;
constexpr int64_t ID_UINT8 = 1125899906844672LL;
constexpr int64_t ID_UINT8B128 = 1125899923621888LL;

}  // namespace mg
}  // namespace aphb200

using namespace aphb200;

extern "C" size_t aph_gptq_marlin_gemm_scratch_bytes(int64_t size_m, int64_t size_n, int64_t size_k) {
  if (size_m <= 0 || size_n <= 0 || size_k <= 0 || size_n % 64 != 0 || size_k % mg::KT != 0) return 0;
  size_t worst = 0;
  for (int nb = 1; nb <= 4; nb *= 2) {
    // upper bound over the slice widths make_plan may pick: grid <= 2 SMs' worth of slots, NT <= 256
    size_t need = (size_t)aphb200::sm_count() * (nb <= 2 ? 2 : 1) * 2 * (8 * nb) * 256 * sizeof(float);
    if (mg::forced_grid_env()) need *= 8;  // development sweeps may use several waves of CTAs
    if (need > worst) worst = need;
  }
  if (size_n % mgtc::NTILE == 0) {
    const size_t need = mg::scratch_bytes_tc(size_m, size_n, size_k);
    if (need > worst) worst = need;
  }
  return worst;
}

extern "C" int aph_gptq_marlin_gemm(void* c, const void* a, const void* b_q_weight, const void* b_scales,
                                    const void* b_zeros, const int32_t* g_idx, const int32_t* perm, int32_t* workspace,
                                    int64_t workspace_ints, int64_t b_q_type, int64_t size_m, int64_t size_n,
                                    int64_t size_k, int64_t lda, int64_t num_groups, int is_k_full, int has_zp,
                                    int use_atomic_add, int use_fp32_reduce, int is_zp_float, int dtype, void* scratch,
                                    size_t scratch_bytes, void* stream) {
  (void)use_atomic_add;   // reduction strategy flags of the reference: results are fp32-reduced here either way
  (void)use_fp32_reduce;
  (void)is_k_full;
  // checks of the reference entry point (gptq_marlin.cu:2271-2440)
  if (has_zp)
    APH_CHECK(b_q_type == mg::ID_UINT4 || b_q_type == mg::ID_UINT8, APH_ERR_INVALID_ARGUMENT,
              "b_q_type must be uint4 or uint8 when has_zp = True. Got = %lld", (long long)b_q_type);
  else
    APH_CHECK(b_q_type == mg::ID_UINT4B8 || b_q_type == mg::ID_UINT8B128, APH_ERR_INVALID_ARGUMENT,
              "b_q_type must be uint4b8 or uint8b128 when has_zp = False. Got = %lld", (long long)b_q_type);
  if (is_zp_float) {
    // HQQ (gptq_marlin.cu:2284-2288,2416-2424): float16 zero points [num_groups, size_n] laid out like the scales
    APH_CHECK(has_zp && b_q_type == mg::ID_UINT4, APH_ERR_INVALID_ARGUMENT, "float zero points need has_zp and uint4 weights");
    APH_CHECK(dtype == APH_DT_F16, APH_ERR_INVALID_ARGUMENT, "Computation type must be float16 (half) when using float zero points.");
    APH_CHECK(num_groups > 1, APH_ERR_INVALID_ARGUMENT, "float zero points need grouped quantisation (num_groups must be != -1)");
    APH_CHECK(size_n % mgtc::NTILE == 0, APH_ERR_UNSUPPORTED,
              "gptq_marlin_gemm: float zero points need size_n divisible by 128 here, got %lld", (long long)size_n);
  }
  const int bits = (b_q_type == mg::ID_UINT8 || b_q_type == mg::ID_UINT8B128) ? 8 : 4;
  // 8-bit weights run on the tensor-memory kernel for every batch size (128-column slices)
  APH_CHECK(bits == 4 || size_n % mgtc::NTILE == 0, APH_ERR_UNSUPPORTED,
            "gptq_marlin_gemm: 8-bit weights need size_n divisible by 128 here, got %lld", (long long)size_n);
  APH_CHECK(perm == nullptr, APH_ERR_INVALID_ARGUMENT,
            "gptq_marlin_gemm: pass activations that are already column-permuted (aph_permute_cols) and perm = NULL");
  APH_CHECK(dtype == APH_DT_F16 || dtype == APH_DT_BF16, APH_ERR_INVALID_ARGUMENT,
            "gpt_marlin_gemm only supports bfloat16 and float16");
  APH_CHECK(size_k % mg::KT == 0, APH_ERR_INVALID_ARGUMENT, "size_k = %lld is not divisible by tile_k = %d",
            (long long)size_k, mg::KT);
  APH_CHECK(size_n % 64 == 0, APH_ERR_INVALID_ARGUMENT, "size_n = %lld is not divisible by tile_n = 64",
            (long long)size_n);
  APH_CHECK(lda % 8 == 0 && aligned16(a), APH_ERR_INVALID_ARGUMENT, "A.stride(0) must be divisible by 8 and A 16-byte aligned");
  APH_CHECK(aligned16(b_q_weight) && aligned16(b_scales), APH_ERR_INVALID_ARGUMENT, "weights / scales must be 16-byte aligned");
  APH_CHECK(num_groups >= 1 && (g_idx != nullptr || size_k % num_groups == 0), APH_ERR_INVALID_ARGUMENT,
            "size_k = %lld is not divisible by the number of groups %lld", (long long)size_k, (long long)num_groups);
  const int64_t group_size = size_k / num_groups;
  APH_CHECK(g_idx != nullptr || num_groups == 1 || group_size == 32 || group_size == 64 || group_size == 128, APH_ERR_INVALID_ARGUMENT,
            "Unsupported group_size = %lld", (long long)group_size);
  if (has_zp) APH_CHECK(b_zeros != nullptr && aligned16(b_zeros), APH_ERR_INVALID_ARGUMENT, "b_zeros missing or misaligned");
  const int64_t min_ws = size_n / 64 * 16;
  APH_CHECK(workspace != nullptr && workspace_ints >= min_ws, APH_ERR_INVALID_ARGUMENT,
            "workspace.numel = %lld is below min_workspace_size = %lld", (long long)workspace_ints, (long long)min_ws);
  if (size_m == 0) return APH_OK;
  if (g_idx != nullptr) {
    // act-order rows whose groups do not follow the k order (is_k_full = False): per-weight scales, marlin_actorder.cuh
    APH_CHECK(!has_zp && b_q_type == mg::ID_UINT4B8, APH_ERR_UNSUPPORTED, "gptq_marlin_gemm: act-order is implemented for uint4b8 weights");
    APH_CHECK(num_groups > 1, APH_ERR_INVALID_ARGUMENT, "For act_order, num_groups must be > 1");
    APH_CHECK(size_m <= 65535 * (int64_t)mgao::MP, APH_ERR_UNSUPPORTED, "gptq_marlin_gemm: size_m too large for the act-order path");
    if (dtype == APH_DT_F16)
      return mgao::launch<__half>(c, a, b_q_weight, b_scales, g_idx, size_m, size_n, size_k, lda, (cudaStream_t)stream);
    return mgao::launch<__nv_bfloat16>(c, a, b_q_weight, b_scales, g_idx, size_m, size_n, size_k, lda, (cudaStream_t)stream);
  }

  mg::Params p;
  p.trace = nullptr;
  p.s = (const uint8_t*)b_scales;
  p.zp = has_zp ? (const uint8_t*)b_zeros : nullptr;
  p.c = c;
  p.counters = workspace;
  p.N = (int)size_n;
  p.K = (int)size_k;
  p.group_size = (int)group_size;
  {
    static const int dbg = mg::env_int("APH_GEMM_DBG");  // development: 1 = skip the math, 2 = skip the loads
    p.dbg = dbg;
  }
  mg::Operands op{a, b_q_weight, b_scales, has_zp ? b_zeros : nullptr, size_m, lda, num_groups, dtype == APH_DT_F16};
  cudaStream_t s = (cudaStream_t)stream;
  if (is_zp_float) return mg::run_typed_tc<__half>(p, op, true, 0 /* HQQ */, (float*)scratch, scratch_bytes, s);
  if (bits == 8 || mg::use_tc_path(size_m, size_n)) {
    if (dtype == APH_DT_F16) return mg::run_typed_tc<__half>(p, op, has_zp != 0, bits, (float*)scratch, scratch_bytes, s);
    return mg::run_typed_tc<__nv_bfloat16>(p, op, has_zp != 0, bits, (float*)scratch, scratch_bytes, s);
  }
  if (dtype == APH_DT_F16) return mg::run_typed<__half>(p, op, has_zp != 0, (float*)scratch, scratch_bytes, s);
  return mg::run_typed<__nv_bfloat16>(p, op, has_zp != 0, (float*)scratch, scratch_bytes, s);
}
