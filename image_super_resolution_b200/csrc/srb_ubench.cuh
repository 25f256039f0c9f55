// Micro-benchmark of tcgen05.mma issue/throughput on this part: cycles per M=128 x N x K=16 MMA as a
// function of N, the operand layout, and how many independent TMEM accumulators are interleaved.
// Measurement tooling only (drives kernel design decisions recorded in DESIGN.md); not on the product path.
#pragma once
#include "srb_ptx.cuh"

namespace srb {

// layout: 0 = SWIZZLE_NONE with the conv kernel's halo-tile strides (LBO=2880, SBO=160)
//         1 = SWIZZLE_NONE dense (LBO=2048 [128 rows x 16 B], SBO=128)
//         2 = SWIZZLE_128B canonical (SBO=1024)
template <int N>
__global__ void __launch_bounds__(128, 1) ubench_mma_kernel(int layout, int racc, int iters, int m64, long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar[4];
    __shared__ uint32_t s_tmem;
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc(&s_tmem, 512);
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = s_tmem;
    const int nwarps = m64 > 0 ? m64 : 1;   // (re-purposed knob) number of issuing warps
    const int wid = threadIdx.x >> 5;
    if (wid < nwarps) {   // warp-uniform control flow; one elected lane issues (no divergent-issue waterfall)
        const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 64 * 1024);
        uint64_t adesc, bdesc;
        if (layout == 0) { adesc = umma_desc_nosw(a_addr, 2880, 160); bdesc = umma_desc_nosw(b_addr, N * 16, 128); }
        else if (layout == 1) { adesc = umma_desc_nosw(a_addr, 2048, 128); bdesc = umma_desc_nosw(b_addr, N * 16, 128); }
        else if (layout == 3) { adesc = umma_desc_nosw(a_addr, 160, 1280); bdesc = umma_desc_nosw(b_addr, N * 16, 128); }   // resblock x row
        else if (layout == 4) { adesc = umma_desc_nosw(a_addr, 160, 640); bdesc = umma_desc_nosw(b_addr, N * 16, 128); }    // resblock hidden row
        else if (layout == 5) { adesc = umma_desc_nosw(a_addr, 176, 1408); bdesc = umma_desc_nosw(b_addr, N * 16, 128); }   // 11-px boxes (odd 16-B pitch)
        else {
            adesc = umma_desc_nosw(a_addr, 16, 1024) | (2ull << 61);
            bdesc = umma_desc_nosw(b_addr, 16, 1024) | (2ull << 61);
        }
        const uint32_t idesc = umma_idesc_f16(128, N, 0);
        const uint32_t tb = tmem_base + wid * racc * N;
        uint64_t* mybar = &bar[wid];
        if (elect_one()) {
            for (int r = 0; r < racc; ++r) umma_f16_ss(tb + r * N, adesc, bdesc, idesc, 0);
            umma_commit(mybar);
        }
        __syncwarp();
        mbar_wait(mybar, 0, 100);
        long long t0 = clock64();
        for (int i = 0; i < iters; i += 4) {
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    for (int r = 0; r < racc; ++r)
                        umma_f16_ss(tb + r * N, adesc + (uint64_t)(k * 2), bdesc, idesc, 1);
            }
            __syncwarp();
        }
        long long t1 = clock64();
        if (elect_one()) umma_commit(mybar);
        __syncwarp();
        mbar_wait(mybar, 1, 101);
        long long t2 = clock64();
        if (blockIdx.x == 0 && threadIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------------------------
// Mixed issue streams (the fused residual block's issuers): warp w issues, per "row", cnt[w] MMAs of N = n[w] round-robin
// over acc[w] accumulators (acc = 1: one dependent chain), then `commits` commits; `rows` rows back to back.
// PAIR: a cluster of two CTAs, cta_group::2 MMAs of M = 256 issued by the leader (each CTA reads its own A rows and
// half of B from its own shared memory).
// ------------------------------------------------------------------------------------------------------------------
struct MixArgs { int nwarps, rows, commits; int n[4], cnt[4], acc[4]; };

template <bool PAIR>
__global__ void __launch_bounds__(128, 1) ubench_mix_kernel(MixArgs a, long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    __shared__ uint64_t bar[8];
    __shared__ uint32_t s_tmem;
    __shared__ long long s_t0;
    uint32_t rank = 0;
    if (PAIR) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    for (int i = threadIdx.x; i < 128 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(&bar[i], 1); fence_mbar_init(); }
    if (threadIdx.x < 32) {
        if (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else tmem_alloc(&s_tmem, 512);
    }
    fence_proxy_async_smem();
    tc_fence_before();
    if (PAIR) { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = s_tmem;
    const int wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_t0 = clock64();
    __syncthreads();
    const long long t0 = s_t0;
    if (wid < a.nwarps && rank == 0) {
        int cb = 0;
        for (int v = 0; v < wid; ++v) cb += a.acc[v] * a.n[v];
        const int n = a.n[wid], cnt = a.cnt[wid], acc = a.acc[wid];
        const uint64_t adesc = umma_desc_nosw(smem_u32(smem) + wid * 8192, 160, 1280);           // x-row boxes
        const uint64_t bdesc = umma_desc_nosw(smem_u32(smem + 64 * 1024), (PAIR ? n / 2 : n) * 16, 128);
        const uint32_t idesc = umma_idesc_f16(PAIR ? 256 : 128, n, 0);
        const uint32_t bstep = (uint32_t)(((PAIR ? n / 2 : n) * 32) >> 4);
        uint64_t* mybar = &bar[wid];
        uint64_t* dummy = &bar[4 + wid];
#pragma unroll 1
        for (int r = 0; r < a.rows; ++r) {
            if (elect_one()) {
                int k = 0;
#pragma unroll 1
                for (int i = 0; i < cnt; ++i) {
                    const uint32_t d = tmem_base + cb + k * n;
                    if (PAIR) {
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                     "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                                     ::"r"(d), "l"(adesc + (uint64_t)(2 * (i & 3))), "l"(bdesc + (uint64_t)((i & 7) * bstep)), "r"(idesc), "r"(1u) : "memory");
                    } else umma_f16_ss(d, adesc + (uint64_t)(2 * (i & 3)), bdesc + (uint64_t)((i & 7) * bstep), idesc, 1u);
                    if (++k == acc) k = 0;
                }
                for (int c = 0; c < a.commits; ++c) {
                    if (PAIR) asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                                           ::"r"(smem_u32(dummy)), "h"((uint16_t)3) : "memory");
                    else umma_commit(dummy);
                }
            }
            __syncwarp();
        }
        const long long t1 = clock64();
        if (elect_one()) {
            if (PAIR) asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                                   ::"r"(smem_u32(mybar)), "h"((uint16_t)3) : "memory");
            else umma_commit(mybar);
        }
        __syncwarp();
        mbar_wait(mybar, 0, 102);
        const long long t2 = clock64();
        if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) { out[1 + wid] = t1 - t0; out[5 + wid] = t2 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = clock64() - t0;
    if (PAIR) { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
    if (threadIdx.x < 32) {
        if (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
        else tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace srb
