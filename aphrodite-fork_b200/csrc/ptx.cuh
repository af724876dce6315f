// Inline-PTX wrappers shared by the sm_100a kernels: mbarrier + 1-D bulk async copy (TMA engine,
// SASS UBLKCP), ldmatrix / movmatrix, mma.sync.m16n8k16, ex2.
#pragma once
#include <cstdint>

namespace aphb200 {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// The suspend-time hint lets a waiting warp sleep inside try_wait until the phase completes (it is woken by the
// barrier) instead of coming back every few hundred cycles: in the tcgen05 GEMM a quarter of all issued instructions
// were such polls (ncu source view, profiles/r02_ncu_source_gemm_tc_gate_up_m64.txt).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(bar),
      "r"(parity), "r"(0x989680)
      : "memory");
}
// 1-D bulk async copy global -> shared, completion counted in bytes on an mbarrier.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

// m16n8k16, A row-major (4 regs), B col-major (2 regs), fp32 accumulate. F16 = true: f16 inputs.
template <bool F16>
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  if constexpr (F16) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  } else {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void ldsm_x2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void lds128(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
// (a & b) | c in one LOP3
__device__ __forceinline__ uint32_t and_or(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ void named_bar_arrive(int id, int nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// 2-D tiled TMA load global -> shared through a tensor map (SASS UTMALDG), completion in bytes on an mbarrier.
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(tmap), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const void* tmap, int c0, int c1, uint32_t bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3}], [%4], %5;" ::"r"(dst),
      "l"(tmap), "r"(c0), "r"(c1), "r"(bar), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void prefetch_tensormap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}

// ---------------------------------------------------------------------------------------------
// tcgen05 (5th-gen tensor core) + tensor memory. Addresses are 32-bit TMEM addresses: lane << 16 | column.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t smem_slot, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_slot), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// all previously issued tcgen05 ops of this thread arrive on the mbarrier when they complete
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]; kind::f16 covers f16 and bf16 inputs, fp32 accumulate
__device__ __forceinline__ void tc_mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// 16 lanes x 8 columns: regs = (lane g, col t), (g+8, t), (g, t+4), (g+8, t+4) with g = laneid/4, t = laneid%4,
// i.e. the register order of an mma.sync m16n8k16 A fragment when a column holds one 16-bit pair.
__device__ __forceinline__ void tmem_st_16x128b_x2(uint32_t taddr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
  asm volatile("tcgen05.st.sync.aligned.16x128b.x2.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
               : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// thread i of the warp reads 16 consecutive columns of lane (base lane + i)
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
// K-major operand tile in shared memory written by TMA with SWIZZLE_128B: rows of 128 bytes, 8-row atoms of 1 KiB.
// (descriptor layout: cute/arch/mma_sm100_desc.hpp SmemDescriptor; version 1 = Blackwell, layout type 2 = SWIZZLE_128B)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}
// instruction descriptor for kind::f16: fp32 accumulate, A and B K-major, fmt 0 = f16, 1 = bf16
__host__ __device__ constexpr uint32_t umma_idesc_f16(int fmt, int m, int n) {
  return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void lds64(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}

// wait with back-off, for single-thread roles (producers, MMA issuer) whose spinning would steal issue slots
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
  uint32_t done;
  for (;;) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    __nanosleep(40);
  }
}
// 16 lanes x 32 columns: four consecutive 16x8 blocks (see tmem_st_16x128b_x2), regs [block][a0..a3]
__device__ __forceinline__ void tmem_st_16x128b_x8(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.16x128b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

// Warp-convergent variants: every lane executes the statement, one elected lane issues the tcgen05 op. Keeping the
// control flow uniform lets ptxas feed the uniform datapath directly instead of looping over "active" lanes.
__device__ __forceinline__ void tc_mma_ts_elect(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_desc_lo, uint32_t b_desc_hi, uint32_t idesc,
                                                uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p, q;\n"
      ".reg .b64 bd;\n"
      "elect.sync _|q, 0xffffffff;\n"
      "setp.ne.b32 p, %5, 0;\n"
      "mov.b64 bd, {%2, %3};\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "r"(b_desc_lo), "r"(b_desc_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_commit_elect(uint32_t bar) {
  asm volatile(
      "{\n"
      ".reg .pred q;\n"
      "elect.sync _|q, 0xffffffff;\n"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
      "}\n" ::"r"(bar)
      : "memory");
}

// Eight K=16 MMAs of one 128-k stage in a single statement: the TMEM column and the descriptor start address
// advance by constants, which keeps the operand arithmetic on the uniform datapath.
__device__ __forceinline__ void tc_mma_ts_stage8(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_desc_lo, uint32_t b_desc_hi, uint32_t idesc,
                                                 uint32_t accumulate_first, uint32_t box_step16) {
  asm volatile(
      "{\n"
      ".reg .pred p, q, t;\n"
      ".reg .b64 bd;\n"
      ".reg .b32 lo, at, lo4;\n"
      "elect.sync _|q, 0xffffffff;\n"
      "setp.ne.b32 p, %5, 0;\n"
      "setp.eq.b32 t, 0, 0;\n"
      "add.u32 lo4, %2, %6;\n"
      "mov.b64 bd, {%2, %3};\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], bd, %4, p;\n"
      "add.u32 lo, %2, 2;\n"
      "mov.b64 bd, {lo, %3};\n"
      "add.u32 at, %1, 8;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [at], bd, %4, t;\n"
      "add.u32 lo, %2, 4;\n"
      "mov.b64 bd, {lo, %3};\n"
      "add.u32 at, %1, 16;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [at], bd, %4, t;\n"
      "add.u32 lo, %2, 6;\n"
      "mov.b64 bd, {lo, %3};\n"
      "add.u32 at, %1, 24;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [at], bd, %4, t;\n"
      "mov.b64 bd, {lo4, %3};\n"
      "add.u32 at, %1, 32;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [at], bd, %4, t;\n"
      "add.u32 lo, lo4, 2;\n"
      "mov.b64 bd, {lo, %3};\n"
      "add.u32 at, %1, 40;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [at], bd, %4, t;\n"
      "add.u32 lo, lo4, 4;\n"
      "mov.b64 bd, {lo, %3};\n"
      "add.u32 at, %1, 48;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [at], bd, %4, t;\n"
      "add.u32 lo, lo4, 6;\n"
      "mov.b64 bd, {lo, %3};\n"
      "add.u32 at, %1, 56;\n"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [at], bd, %4, t;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "r"(b_desc_lo), "r"(b_desc_hi), "r"(idesc), "r"(accumulate_first), "r"(box_step16)
      : "memory");
}
// release / acquire fence at GPU scope (cheaper than the sequentially consistent fence __threadfence() emits)
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// programmatic dependent launch (PDL): wait = all prerequisite grids completed and their writes are visible;
// launch_dependents = the next kernel of the stream may be scheduled once every CTA of this grid issued it or exited
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
}

// ---------------------------------------------------------------------------------------------
// Thread-block clusters: the cluster-wide barrier (one barrier, alternating phases; every thread of every CTA of the
// cluster arrives, exited threads count as arrived) and stores into another CTA's shared memory (DSMEM).
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// Both phases of a two-phase cluster rendezvous for a thread that has nothing to contribute (warp-specialised kernels:
// producer / dequantiser / idle warps). Not .aligned: the lanes of a warp may reach it from different places. Every
// thread of the cluster takes part in every phase, so nothing depends on how exited threads are counted.
__device__ __forceinline__ void cluster_sync_twice_unaligned() {
  asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;\n\tbarrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// address of the same shared-memory location in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t cluster_map_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

// ---------------------------------------------------------------------------------------------
// Position in a ring of DEPTH pipeline slots whose items are consumed by TWO groups taking alternate items
// (item `it` belongs to group it & 1). Each group owns a sub-ring of its own: group 0 the even slots
// (ceil(DEPTH/2) of them), group 1 the odd slots (floor(DEPTH/2)). A slot is therefore only ever filled for one
// group, and every thread that waits on a slot's mbarriers meets EVERY phase of that slot in order.
//
// Why not slot = it % DEPTH: with an odd DEPTH a slot then alternates between the groups, so a group meets a slot
// only every second phase and always waits with the same parity. mbarrier.try_wait.parity(P) is true whenever the
// barrier's current phase has parity != P; it cannot tell "my phase 2k completed" from "phase 2k-2 completed and the
// other group's phase 2k-1 is still in flight". A group that runs ahead of the other group's load (TMA loads complete
// out of order) then reads a slot that still holds an older stage: a wrong result, silently. This was the cause of
// the round-1 "odd ring depth" failures (about 0.1 % of calls); `-DAPH_RING_LEGACY` restores the flat mapping and
// `-DAPH_RING_FAULT` makes the producers land the odd items late, which turns the rare failure into a deterministic
// one (tools/ring_fault_proof.py).
//
// slot(it) / phase(it) describe item `it`; advance(it) must be called once for every item the thread handles, in
// order. A thread may handle all items or only those of one group.
template <int DEPTH>
struct RingCursor {
  static_assert(DEPTH >= 2, "a ring shared by two groups needs a slot per group");
  static constexpr int D0 = (DEPTH + 1) / 2, D1 = DEPTH / 2;
  int pos0 = 0, pos1 = 0;
  uint32_t ph0 = 0, ph1 = 0;
  __device__ __forceinline__ RingCursor() {}
  __device__ __forceinline__ explicit RingCursor(int) {}  // same constructor as GroupCursor (the group is not used)
#ifdef APH_RING_LEGACY
  __device__ __forceinline__ int slot(int it) const { return it % DEPTH; }
  __device__ __forceinline__ uint32_t phase(int it) const { return (uint32_t)(it / DEPTH) & 1u; }
  __device__ __forceinline__ void advance(int) {}
#else
  __device__ __forceinline__ int slot(int it) const { return (it & 1) ? 1 + 2 * pos1 : 2 * pos0; }
  __device__ __forceinline__ uint32_t phase(int it) const { return (it & 1) ? ph1 : ph0; }
  __device__ __forceinline__ void advance(int it) {
    if (it & 1) {
      if (++pos1 == D1) {
        pos1 = 0;
        ph1 ^= 1u;
      }
    } else {
      if (++pos0 == D0) {
        pos0 = 0;
        ph0 ^= 1u;
      }
    }
  }
#endif
};

// The same mapping for a thread that only ever handles the items of ONE group (its & 1 == grp): a plain counter over
// the group's sub-ring. `its` (the item's global index) is only looked at by the APH_RING_LEGACY build.
template <int DEPTH>
struct GroupCursor {
  static_assert(DEPTH >= 2, "a ring shared by two groups needs a slot per group");
  int pos = 0, n, base;
  uint32_t ph = 0;
  __device__ __forceinline__ explicit GroupCursor(int grp) : n(grp ? DEPTH / 2 : (DEPTH + 1) / 2), base(grp) {}
#ifdef APH_RING_LEGACY
  __device__ __forceinline__ int slot(int its) const { return its % DEPTH; }
  __device__ __forceinline__ uint32_t phase(int its) const { return (uint32_t)(its / DEPTH) & 1u; }
  __device__ __forceinline__ void advance(int = 0) {}
#else
  __device__ __forceinline__ int slot(int) const { return base + 2 * pos; }
  __device__ __forceinline__ uint32_t phase(int) const { return ph; }
  __device__ __forceinline__ void advance(int = 0) {
    if (++pos == n) {
      pos = 0;
      ph ^= 1u;
    }
  }
#endif
};

// -DAPH_RING_FAULT (development, tools/ring_fault_proof.py): how long a producer holds back an odd item while the
// even item behind it is already in flight
__device__ __forceinline__ void ring_fault_delay() {
  const long long t0 = clock64();
  while (clock64() - t0 < 12000) {
  }
}

}  // namespace ptx
}  // namespace aphb200
