// bqa_common.cuh — sm_100a device primitives shared by every kernel in libbqa.
//
// Thin inline-PTX wrappers (mbarrier, TMA, tcgen05/TMEM), the per-element dropout hash,
// bf16 packing helpers and warp/block reductions.  Nothing here is a translation of the
// reference (which ships no native code); encodings follow the PTX ISA
// for sm_100a.
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace bqa {

typedef __nv_bfloat16 bf16;

// ----------------------------------------------------------------------------------------
// address helpers
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ uint64_t global_timer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ----------------------------------------------------------------------------------------
// mbarrier
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

// Bounded wait: a protocol bug must surface as a trapped kernel (CUDA error on the host),
// never as a hung GPU.  The bound (4 s of wall time) is far above any legitimate wait.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  // the global timer is slow to read: it is consulted only every 1024 failed polls (each poll already
  // suspends the thread for a hardware-defined time), never on the fast path
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3ffu) == 0) {
      const uint64_t now = global_timer_ns();
      if (t0 == 0) {
        t0 = now;
      } else if (now - t0 > 4000000000ull) {
        printf("bqa: mbarrier timeout block (%d,%d,%d) thread %d bar@%u parity %u\n", blockIdx.x, blockIdx.y,
               blockIdx.z, threadIdx.x, smem_u32(bar), parity);
        __trap();
      }
    }
  }
}

// Read-only global loads that the compiler may not move: used for prefetches that must be ISSUED at a
// given point of a barrier-synchronised loop (a plain load may be sunk to its first use).
__device__ __forceinline__ uint32_t ldg_issue_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
// 4-byte asynchronous global -> shared copy (LDGSTS): the value never occupies a register while in flight
__device__ __forceinline__ void cp_async_4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ float ldg_issue_f32(const float* p) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// ----------------------------------------------------------------------------------------
// Cluster launch control: a running cluster asks the hardware to cancel a cluster of the same grid
// that has not started yet and takes over its work (its blockIdx) — dynamic tile scheduling without
// a software queue.  The 16-byte response lands in shared memory (of every CTA of the cluster when
// `multicast`) and completes 16 bytes on the mbarrier at the same offset.
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ void clc_try_cancel(void* response, uint64_t* bar, bool multicast) {
  if (multicast)
    asm volatile(
        "clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.multicast::cluster::all.b128 "
        "[%0], [%1];" ::"r"(smem_u32(response)), "r"(smem_u32(bar))
        : "memory");
  else
    asm volatile("clusterlaunchcontrol.try_cancel.async.shared::cta.mbarrier::complete_tx::bytes.b128 [%0], [%1];" ::"r"(
                     smem_u32(response)),
                 "r"(smem_u32(bar))
                 : "memory");
}
// true + blockIdx.x of the first CTA of the cancelled cluster, or false when no cluster was left
__device__ __forceinline__ bool clc_decode(const void* response, uint32_t& ctaid_x) {
  uint32_t valid, x, y, z;
  asm volatile(
      "{\n\t"
      ".reg .pred p1;\n\t"
      ".reg .b128 r;\n\t"
      "ld.shared.b128 r, [%4];\n\t"
      "clusterlaunchcontrol.query_cancel.is_canceled.pred.b128 p1, r;\n\t"
      "selp.u32 %3, 1, 0, p1;\n\t"
      "mov.u32 %0, 0; mov.u32 %1, 0; mov.u32 %2, 0;\n\t"
      "@p1 clusterlaunchcontrol.query_cancel.get_first_ctaid.v4.b32.b128 {%0, %1, %2, _}, r;\n\t"
      "}"
      : "=r"(x), "=r"(y), "=r"(z), "=r"(valid)
      : "r"(smem_u32(response))
      : "memory");
  ctaid_x = x;
  return valid != 0;
}

// Programmatic dependent launch: a kernel launched with the programmatic-stream-serialization attribute may
// start while its predecessor in the stream is still running; everything it does before pdl_wait() (barrier
// init, TMEM allocation, descriptor prefetch) overlaps the predecessor's tail.  pdl_wait() returns once the
// predecessor has completed and its writes are visible; pdl_trigger() lets the successor start as soon as
// resources free up.  Both are no-ops for a normally launched kernel.  Nothing the predecessor may have
// written is touched before pdl_wait(), and the trigger comes after the wait, so completion is transitive.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// One lane of a converged warp (the others get false).  Code that issues tcgen05.mma / TMA from inside
// `if (elect_one())` within warp-uniform control flow keeps descriptors in uniform registers; issuing from
// a `threadIdx.x == k` branch instead makes the compiler wrap every instruction in an election loop.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}
// warp index as a value the compiler knows to be warp-uniform
__device__ __forceinline__ int uniform_warp_idx() { return __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0); }

// generic-proxy writes to shared memory -> visible to the async proxy (TMA store, UMMA reads)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ----------------------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor)
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, int32_t c0, int32_t c1,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1), "r"(smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, int32_t c0, int32_t c1,
                                            int32_t c2, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1), "r"(c2),
      "r"(smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ void tma_store_2d(const CUtensorMap* m, const void* smem_src, int32_t c0, int32_t c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}

__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, const void* smem_src, int32_t c0, int32_t c1,
                                             int32_t c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* m, const void* smem_src, int32_t c0,
                                                  int32_t c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(m)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}

__device__ __forceinline__ void tma_commit_group() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }

template <int N>
__device__ __forceinline__ void tma_wait_group_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

template <int N>
__device__ __forceinline__ void tma_wait_group() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ----------------------------------------------------------------------------------------
// tcgen05 / TMEM
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "r"(ncols)
               : "memory");
}

__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate; one thread issues.
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// mbarrier arrives once every tcgen05.mma issued so far by this thread has completed
// (implies tcgen05.fence::before_thread_sync).
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives TMEM lane
// (lane field of taddr, which must be 32*(warp_id%4)) + i, columns [col, col+32).
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}

__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (128 rows x 16 bf16 per instruction) is read from tensor
// memory, lane = row, two consecutive K elements per 32-bit column (element k of a row in column k / 2, low half
// first).  A must be K-major in this form.
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// registers -> TMEM, mirror images of tmem_ld_32x32 / tmem_ld_32x16
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
      "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
      "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]),
      "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]),
      "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}

__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ----------------------------------------------------------------------------------------
// CTA-pair (cta_group::2) variants: two CTAs of a cluster, on the two SMs of one TPC, execute
// one 256-row MMA; the even-ranked CTA issues it and owns the pipeline barriers.
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
}

// arrive on the barrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t cta) {
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t"
      "}"
      ::"r"(smem_u32(bar)), "r"(cta)
      : "memory");
}

// TMA load whose completion bytes are credited to the barrier of the pair's even CTA
// (bit 24 of a shared::cluster address selects the CTA within the pair)
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, int32_t c0, int32_t c1,
                                                 uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(c0), "r"(c1),
      "r"(smem_u32(bar) & 0xFEFFFFFFu)
      : "memory");
}

__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_result, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "r"(ncols)
               : "memory");
}

__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                               uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

// commit that arrives on the barrier at this offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
      : "memory");
}

// ----------------------------------------------------------------------------------------
// UMMA descriptors (bf16, 128-byte swizzle, BLOCK_K = 64 elements)
// ----------------------------------------------------------------------------------------
// Shared-memory matrix descriptor, sm_100 format:
//   [0,14)  start address >> 4      [16,30) leading byte offset >> 4
//   [32,46) stride byte offset >> 4 [46,48) version = 1      [61,64) layout (2 = SWIZZLE_128B)
__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}

// K-major operand tile: rows x 64 bf16 (one 128-byte swizzle span per row), 8-row atoms of 1024 B.
__device__ __forceinline__ uint64_t umma_desc_kmajor(uint32_t smem_addr) { return umma_smem_desc(smem_addr, 0, 1024); }
// MN-major operand tile: [mn_atom][k rows][64 bf16]; 8 k-rows per 1024-byte atom, mn atoms
// `atom_stride_bytes` apart (= k rows per tile * 128 B).
__device__ __forceinline__ uint64_t umma_desc_mnmajor(uint32_t smem_addr, uint32_t atom_stride_bytes) {
  return umma_smem_desc(smem_addr, atom_stride_bytes, 1024);
}

// Instruction descriptor for kind::f16, bf16 x bf16 -> fp32.
//   [4,6) D format (1 = f32)  [7,10) A format (1 = bf16)  [10,13) B format (1 = bf16)
//   [15] A major (1 = MN)     [16] B major (1 = MN)       [17,23) N >> 3   [24,29) M >> 4
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int m, int n, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(a_mn_major) << 15) |
         (static_cast<uint32_t>(b_mn_major) << 16) | (static_cast<uint32_t>(n >> 3) << 17) |
         (static_cast<uint32_t>(m >> 4) << 24);
}

// 16-byte vector reduction into global memory (sm_90+): four fp32 adds in one instruction
__device__ __forceinline__ void red_add_f32x4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// ----------------------------------------------------------------------------------------
// numerics helpers
// ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) {
  __nv_bfloat162 t = *reinterpret_cast<__nv_bfloat162*>(&v);
  return __bfloat1622float2(t);
}

// GELU (erf form: transformers activations.py -> F.gelu) and its derivative from ONE shared core.
//   Phi(x) = 0.5 (1 + erf(x / sqrt 2)),  gelu(x) = x Phi(x),  gelu'(x) = Phi(x) + x phi(x)
// erf(z), z >= 0, by Abramowitz-Stegun 7.1.25: 1 - (a1 t + a2 t^2 + a3 t^3) exp(-z^2),
// t = 1 / (1 + p z), |error| <= 2.5e-5 — two orders below the bf16 resolution of the stored
// results — using one MUFU.RCP, one MUFU.EX2 and no branches, so that the GEMM epilogue that
// applies it keeps pace with the tensor cores.  exp(-z^2) = exp(-x^2 / 2) is also the Gaussian
// the derivative needs.
__device__ __forceinline__ void gelu_and_grad(float x, float& g, float& dg) {
  // w = |x| * sqrt(log2(e) / 2): exp(-x^2/2) = 2^(-w^2);  p z = (p / sqrt(log2 e)) w
  const float w = fabsf(x) * 0.84932180028801904f;
  float e, t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-w * w));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.39169197f, w, 1.0f)));
  // h = 0.5 * (1 - erf(z)) = 0.5 t (a1 + a2 t + a3 t^2) e
  float poly = fmaf(t, 0.3739278f, -0.0479399f);
  poly = fmaf(t, poly, 0.1740121f);
  const float h = poly * t * e;
  const float phi_cdf = 0.5f + copysignf(0.5f - h, x);
  g = x * phi_cdf;
  dg = fmaf(x * 0.39894228040143268f, e, phi_cdf);
}

// ---- packed fp32 pairs (FFMA2 / FMUL2 / FADD2 on sm_100): one issue slot for two lanes of fp32 work.  The epilogues
// and row kernels here are bound by instruction issue, not by the fp32 pipe, so halving the issue count of their
// multiply-adds is worth the explicit pairing.  A pair lives in a 64-bit register (lo = first element).
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t f2_add(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t f2_splat(float v) { return f2_pack(v, v); }

// gelu_and_grad for two elements at once: the same operations in the same order as the scalar version (bit-identical
// results), with the multiply-adds issued as packed pairs.
__device__ __forceinline__ void gelu_and_grad2(float x0, float x1, float& g0, float& g1, float& dg0, float& dg1) {
  const uint64_t x = f2_pack(x0, x1);
  const uint64_t ax = f2_pack(fabsf(x0), fabsf(x1));
  const uint64_t w = f2_mul(ax, f2_splat(0.84932180028801904f));
  const uint64_t nw = f2_mul(ax, f2_splat(-0.84932180028801904f));
  float a0, a1, u0, u1, e0, e1, t0, t1;
  f2_unpack(f2_mul(nw, w), a0, a1);
  f2_unpack(f2_fma(f2_splat(0.39169197f), w, f2_splat(1.0f)), u0, u1);
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(a0));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(a1));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(u0));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(u1));
  const uint64_t t = f2_pack(t0, t1), e = f2_pack(e0, e1);
  uint64_t poly = f2_fma(t, f2_splat(0.3739278f), f2_splat(-0.0479399f));
  poly = f2_fma(t, poly, f2_splat(0.1740121f));
  const uint64_t h = f2_mul(f2_mul(poly, t), e);
  float d0, d1;
  f2_unpack(f2_fma(h, f2_splat(-1.0f), f2_splat(0.5f)), d0, d1);  // 0.5 - h, one rounding like the scalar subtraction
  const uint64_t phi = f2_add(f2_splat(0.5f), f2_pack(copysignf(d0, x0), copysignf(d1, x1)));
  f2_unpack(f2_mul(x, phi), g0, g1);
  f2_unpack(f2_fma(f2_mul(x, f2_splat(0.39894228040143268f)), e, phi), dg0, dg1);
}

__device__ __forceinline__ float gelu_erf(float x) {
  float g, dg;
  gelu_and_grad(x, g, dg);
  return g;
}

__device__ __forceinline__ float gelu_erf_grad(float x) {
  float g, dg;
  gelu_and_grad(x, g, dg);
  return dg;
}

// Stateless dropout: a Weyl step plus one round of a 32-bit avalanche mixer keyed by (k0, k1),
// which the host derives from (seed, step, site).  One hash serves TWO consecutive elements (its low and high
// 16 bits), decision = (16-bit value >= thresh).  The same (key, element index) gives the same
// decision in the forward kernel and in every backward kernel that needs the mask.
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352dU;
  x ^= x >> 15;
  x *= 0x846ca68bU;
  x ^= x >> 16;
  return x;
}

struct DropKey {
  uint32_t k0, k1;
  uint32_t thresh;  // drop iff 16-bit value < thresh ; thresh = round(p * 65536); 0 disables
  float scale;      // 1 / (1 - p)
};

__device__ __forceinline__ uint32_t drop_hash(const DropKey& dk, uint32_t pair_idx) {
  return mix32(pair_idx * 0x9E3779B1u + dk.k0) ^ dk.k1;
}

__device__ __forceinline__ bool drop_keep(const DropKey& dk, uint32_t idx) {
  const uint32_t h = drop_hash(dk, idx >> 1);
  return ((idx & 1u) ? (h >> 16) : (h & 0xffffu)) >= dk.thresh;
}

// decisions for elements idx_even and idx_even + 1 (idx_even must be even)
__device__ __forceinline__ void drop_keep2(const DropKey& dk, uint32_t idx_even, bool& keep0, bool& keep1) {
  const uint32_t h = drop_hash(dk, idx_even >> 1);
  keep0 = (h & 0xffffu) >= dk.thresh;
  keep1 = (h >> 16) >= dk.thresh;
}

// keep decisions for the 32 consecutive elements starting at `idx`, one bit each
__device__ __forceinline__ uint32_t keep_bits32(const DropKey& dk, uint32_t idx) {
  uint32_t bits = 0;
  if ((idx & 1u) == 0u) {
#pragma unroll
    for (int e = 0; e < 32; e += 2) {
      const uint32_t h = drop_hash(dk, (idx + e) >> 1);
      bits |= ((h & 0xffffu) >= dk.thresh ? 1u : 0u) << e;
      bits |= ((h >> 16) >= dk.thresh ? 1u : 0u) << (e + 1);
    }
  } else {
#pragma unroll
    for (int e = 0; e < 32; ++e) bits |= (drop_keep(dk, idx + e) ? 1u : 0u) << e;
  }
  return bits;
}

// three-input maximum (one FMNMX3 on sm_100)
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace bqa
