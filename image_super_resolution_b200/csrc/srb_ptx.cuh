// sm_100a PTX wrappers used by the conv kernels: mbarrier, TMA (cp.async.bulk[.tensor]),
// tcgen05 (TMEM alloc / mma / commit / ld) and UMMA descriptor builders.
// Hand-written for B200; nothing here is taken from the reference (which has no native code).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

// Probe hooks (bottleneck probes, clock traces) exist only in the -DSRB_DEBUG build that tools/ loads
// (libsrb200_dbg.so); in the product library every such branch is a compile-time constant and disappears.
#ifdef SRB_DEBUG
#define SRB_DBG(expr) (expr)
#else
#define SRB_DBG(expr) (0)
#endif

namespace srb {

// ------------------------------------------------------------------------------------------
// Bounded waits: a protocol bug must not hang the GPU box (that is a strike); after a generous
// spin budget the kernel records where it was stuck, raises the abort flag and the launch drains.
// The flag is mirrored into host-mapped pinned memory, so the host can poll it without a device
// synchronisation (srb_poll_abort: every sync point of the Python layer raises on it).
// ------------------------------------------------------------------------------------------
__device__ unsigned int g_srb_timeout_info[4];
__device__ volatile unsigned int g_srb_abort;   // set by the first wait that runs out of budget; every later wait bails out
__device__ unsigned int* g_srb_abort_host;      // host-mapped mirror: [0] flag, [1..4] timeout info (set up by srb_create)
#ifndef SRB_WAIT_BUDGET_LOG2
#define SRB_WAIT_BUDGET_LOG2 32                 // 2^32 cycles: ~2.3 s at 1.9 GHz (time-sliced / profiled runs stay far below)
#endif

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
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
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// try_wait that lets the hardware park the thread for up to `ns` nanoseconds (woken when the phase completes) instead
// of returning at once: a parked warp issues nothing, a polling one burns issue slots, shared-memory reads and power
__device__ __forceinline__ bool mbar_try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t ns) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(ns)
        : "memory");
    return ok != 0;
}
// tag identifies the wait site.  On a protocol bug the kernel does not hang or trap: the first wait to
// exceed its budget records where it was, raises g_srb_abort, and every wait returns at once so the
// launch drains (with garbage results); the host reads the flag (srb_check_device) and reports it.
// SLEEP_NS > 0: back off between polls — for consumer warps that share a scheduler with a latency-critical issuing
// warp (a spinning warp steals the issuer's issue slots).
template <int SLEEP_NS = 0>
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, uint32_t tag) {
    if (mbar_try_wait(bar, parity)) return;
    long long t0 = clock64();
    for (;;) {
        // the abort flag lives in global memory (an L2 round trip per look): poll the barrier many times per look, or
        // every hand-off in the pipeline pays that latency
#pragma unroll 1
        for (int i = 0; i < 64; ++i) {
            if constexpr (SLEEP_NS < 0) {            // parked by the hardware for up to -SLEEP_NS ns per try
                if (mbar_try_wait_hint(bar, parity, (uint32_t)(-SLEEP_NS))) return;
            } else {
                if (mbar_try_wait(bar, parity)) return;
                if (SLEEP_NS > 0) __nanosleep(SLEEP_NS);
            }
        }
        if (g_srb_abort) return;
        if (clock64() - t0 > (1ll << SRB_WAIT_BUDGET_LOG2)) {
            if (atomicExch((unsigned int*)&g_srb_abort, 1u) == 0u) {
                g_srb_timeout_info[0] = 0xdead0000u | tag;
                g_srb_timeout_info[1] = blockIdx.x;
                g_srb_timeout_info[2] = threadIdx.x;
                g_srb_timeout_info[3] = parity;
                unsigned int* h = g_srb_abort_host;
                if (h != nullptr) {
                    h[1] = 0xdead0000u | tag; h[2] = blockIdx.x; h[3] = threadIdx.x; h[4] = parity;
                    __threadfence_system();
                    *reinterpret_cast<volatile unsigned int*>(h) = 1u;
                }
            }
            __threadfence_system();
            return;
        }
    }
}

// ------------------------------------------------------------------------------------------
// TMA
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// 4-D tiled tensor load: coordinates innermost first
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// 5-D tiled tensor load (a whole row of 8-px group boxes in one instruction: srb_tail.cuh)
__device__ __forceinline__ void tma_load_5d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, int c2,
                                            int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}
// 1-D bulk copy global -> shared (bytes % 16 == 0, 16-B aligned both sides)
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// 4-D tiled tensor store shared -> global (bulk async group); out-of-bounds parts of the box are clipped
__device__ __forceinline__ void tma_store_4d(const void* tmap, const void* smem_src, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
        ::"l"(tmap), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed bulk stores of this thread have finished READING shared memory (the source may be overwritten)
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// Programmatic dependent launch: a kernel launched with programmaticStreamSerialization may start while its predecessor
// in the stream is still running; pdl_wait() blocks until the predecessor has completed and its writes are visible (a
// no-op for a normal launch), pdl_launch_dependents() lets the successor's CTAs take the SMs this grid's CTAs vacate.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// named barrier among `count` threads (id 1..15; id 0 is __syncthreads)
__device__ __forceinline__ void named_bar_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// ------------------------------------------------------------------------------------------
// tcgen05 / TMEM
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {  // whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {  // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T ; one thread issues for the CTA
__device__ __forceinline__ void umma_f16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on an mbarrier once all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 32-bit, 8 consecutive columns per thread (thread i of the warp <-> TMEM lane base+i)
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}

// ------------------------------------------------------------------------------------------
// UMMA descriptors (K-major, SWIZZLE_NONE "interleaved" canonical layout):
//   in 16-byte units  ((8, m), 2) : ((1, SBO), LBO)
//   - a core matrix is 8 rows x 16 B, rows 16 B apart (128 contiguous bytes)
//   - SBO = byte distance between core matrices that are adjacent along M (or N)
//   - LBO = byte distance between the two core matrices along K of one K=16 instruction
// bits: [0,14) addr>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout=0
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t umma_desc_nosw(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// kind::f16 instruction descriptor: fp32 accumulate, A/B K-major. fmt: 0 = fp16, 1 = bf16
__host__ __device__ constexpr uint32_t umma_idesc_f16(int M, int N, int fmt) {
    return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}

}  // namespace srb

// ------------------------------------------------------------------------------------------
// CTA pairs (thread-block cluster of two CTAs on one TPC, tcgen05 cta_group::2): one M = 256 MMA issued by the leader CTA
// works on both CTAs' A rows / TMEM accumulators and reads HALF of B from each CTA's shared memory.
// ------------------------------------------------------------------------------------------
namespace srb {

__device__ __forceinline__ void prefetch_l2(const void* p) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {       // every thread of every CTA of the cluster
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
    return r;
}
// arrive on a barrier given by a shared::cluster address (this CTA's or the peer's).  Default semantics (release at CTA
// scope): what is handed over lives in the ARRIVING CTA's own shared memory / TMEM and is consumed by that SM's tensor core;
// the leader only observes the barrier and issues the MMA.  (Measured: `.release.cluster` makes every arrival a cluster-scope
// memory barrier — 1 500 to 3 000 clk per hand-off in the epilogues, profiles/rb_trace_pair_r2.log.)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {   // acquire at cluster scope
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// bounded wait like mbar_wait, observing arrivals from the peer CTA
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity, uint32_t tag) {
    if (mbar_try_wait_cluster(bar, parity)) return;
    long long t0 = clock64();
    for (;;) {
#pragma unroll 1
        for (int i = 0; i < 64; ++i)
            if (mbar_try_wait_cluster(bar, parity)) return;
        if (g_srb_abort) return;
        if (clock64() - t0 > (1ll << SRB_WAIT_BUDGET_LOG2)) {
            if (atomicExch((unsigned int*)&g_srb_abort, 1u) == 0u) {
                g_srb_timeout_info[0] = 0xdead0000u | tag;
                g_srb_timeout_info[1] = blockIdx.x;
                g_srb_timeout_info[2] = threadIdx.x;
                g_srb_timeout_info[3] = parity;
                unsigned int* h = g_srb_abort_host;
                if (h != nullptr) {
                    h[1] = 0xdead0000u | tag; h[2] = blockIdx.x; h[3] = threadIdx.x; h[4] = parity;
                    __threadfence_system();
                    *reinterpret_cast<volatile unsigned int*>(h) = 1u;
                }
            }
            __threadfence_system();
            return;
        }
    }
}
// 4-D tiled tensor load into THIS CTA's shared memory whose completion is signalled on a barrier of either CTA of the pair
// (`bar_cluster_addr`: shared::cluster address, e.g. the leader's barrier from mapa_u32)
__device__ __forceinline__ void tma_load_4d_pair(void* smem_dst, const void* tmap, uint32_t bar_cluster_addr, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
        "[%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_result, uint32_t ncols) {  // whole warp, in both CTAs
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem of both CTAs] += A[each CTA's 128 rows] * B[N/2 rows from each CTA]^T, M = 256; issued by one thread of the leader CTA
__device__ __forceinline__ void umma_f16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs once all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

}  // namespace srb
