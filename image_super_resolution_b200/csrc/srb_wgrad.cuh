// Weight gradient of a 3x3 conv on tcgen05/TMEM:  dW[co][ci][dy][dx] = sum_p dY[p][co] * X[p + (dy-1, dx-1)][ci]
// (autograd of aten::_convolution reached from train.py:57).
//
// GEMM view: reduction K = pixels, both operands are C8-plane tiles read "MN-major" (a 16-byte granule = 8 channels of
// one pixel, consecutive pixels 16 B apart: 8 pixels x 8 channels = one 128-byte core matrix, LBO = 128 B = next 8 pixels,
// SBO = next 8 channels).  The THREE dy TAPS ARE FOLDED INTO M and the input channels are N:
//     D[(s, co)][ci] += A[(s, co)][p] * B[ci][p],     A = dY rows r+s (s = 2-dy), B = X row r shifted by dx pixels,
// M = 3 x 32 output channels (+ 32 idle lanes), N = the unit's input channels (up to 160, exact: no padding to 128),
// K = 16 pixels (one tile row) per MMA, one TMEM accumulator per dx tap persisting over all pixel tiles of the CTA.
// The fold needs no copies: the dY tile (16 px x 10 rows x 4 planes, y halo) is loaded by a tensor map whose dimensions
// are ordered (pixel, plane, row), so shared memory holds [row][plane][16 px] and the 8-channel groups (s, plane) of
// rows r, r+1, r+2 are 256 B apart — a uniform SBO.  The dx tap is a 16-byte shift of the X tile's start address
// (18 px x 8 rows x N/8 planes, x halo).  Against one N = 32 MMA per tap (40 % of the tensor rate, input channels padded
// to 128) this is 3 MMAs of N = cin per tile row instead of 9 (or 18).  A unit = (conv, <= 160 input channels, 32 output
// channels); splitk CTAs share its pixel tiles and add their partial dW to global fp32 with red.global.add.
#pragma once
#include <cuda.h>
#include "srb_ptx.cuh"

namespace srb {

struct WgradUnit {
    int tm;                       // batch: index of the conv's tensor-map pair: tmaps[2*tm] = X, tmaps[2*tm+1] = dY
    int x_c8_off, dy_c8_off;      // first plane of the cin block / cout block
    int xplanes;                  // planes in the X box (even): the MMA's N = 8 * xplanes
    int cin_blk, cout_blk, cin_total, cin0, cout0;   // real channels of the block; dW indexing [cout][cin_total][3][3]
    int N, H, W, tiles_x, tiles_y, num_tiles;
    float scale;
    float* dw;
    float* db;                    // bias gradient of the cout block (this unit sums the dY tiles it loads anyway), or nullptr
};

// KS = 3: the 3x3 convs (4 dY planes = 32 output channels per unit, N up to 160 input channels).
// KS = 9: the 9x9 head / tail convs (utils/models.py:625, :636) in the same mapping — one dY plane (8 output channels, the
// tail's 3 + zeros) x 9 row shifts = 9 of the 16 M groups, 9 dx accumulators of N <= 32 input channels (the head's 3 + zeros).
template <int KS>
struct WgradGeomT {
    static constexpr int TW = 16, TH = 8, PAD = KS / 2;
    static constexpr int XW = TW + 2 * PAD, DYPLANES = (KS == 3) ? 4 : 1, DYH = TH + 2 * PAD;
    static constexpr int MAXPLANES = (KS == 3) ? 20 : 4;                      // KS accumulators x 8 * MAXPLANES columns <= 512
    static constexpr int XROWP = XW * 16, XPLANE = TH * XROWP;                // 3x3: 288, 2304;  9x9: 384, 3072
    static constexpr int X_STRIDE = 46080;                                    // = 20 planes of the 3x3 geometry
    static constexpr int DYGRP = TW * 16, DYROWP = DYPLANES * DYGRP, DYBYTES = DYH * DYROWP;   // 3x3: 256, 1024, 10240;  9x9: 256, 256, 4096
    static constexpr int DY_STRIDE = 10240;
    static constexpr int STAGE_STRIDE = X_STRIDE + DY_STRIDE;                 // 56320
    static constexpr int STAGES = 4;
    static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_STRIDE + 1024 + 256;   // the idle M lanes read past a dY tile (<= 1 KB past the stage)
    static constexpr int NOUT = 8 * DYPLANES;
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
    static_assert(MAXPLANES * XPLANE <= X_STRIDE && DYBYTES <= DY_STRIDE, "stage layout");
    static_assert((TH - 1) * DYROWP + 16 * DYGRP <= DY_STRIDE + 1024, "A operand over-read stays inside the allocation");
    static_assert(KS * MAXPLANES * 8 <= 512, "TMEM columns");
};
using WgradGeom = WgradGeomT<3>;

// kind::f16 instruction descriptor with both operands MN-major
__host__ __device__ constexpr uint32_t umma_idesc_f16_mn(int M, int N, int fmt) {
    return umma_idesc_f16(M, N, fmt) | (1u << 15) | (1u << 16);
}

template <bool BF16, int KS>
__device__ __forceinline__ void wgrad_body(const CUtensorMap* tmap_x, const CUtensorMap* tmap_dy, const WgradUnit& u, int part,
                                           int splitk, int dbg_skip) {
    using G = WgradGeomT<KS>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* s_st = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::STAGES * G::STAGE_STRIDE + 1024);
    uint64_t* full = bars;          // [4]
    uint64_t* empty = bars + 4;     // [4]
    uint64_t* done = bars + 8;      // [1]
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 10);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        // a stage is released by the MMA commit and, when the unit also owns a bias gradient, by the two summing warps
        for (int s = 0; s < G::STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], u.db != nullptr ? 3 : 1); }
        mbar_init(done, 1);
        fence_mbar_init();
        tma_prefetch_desc(tmap_x);
        tma_prefetch_desc(tmap_dy);
    }
    if (warp == 0) { __syncwarp(); tmem_alloc(s_tmem, 512); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    const int tiles_per_img = u.tiles_x * u.tiles_y;
    const int NN = u.xplanes * 8;                         // MMA N

    if (warp == 0) {
        const uint32_t bytes = (uint32_t)(u.xplanes * G::XPLANE + G::DYBYTES);
        int s = 0; uint32_t ph = 0;
        for (int tile = part; tile < u.num_tiles; tile += splitk) {
            const int n = tile / tiles_per_img;
            const int r = tile - n * tiles_per_img;
            const int ty = r / u.tiles_x, tx = r - ty * u.tiles_x;
            mbar_wait(&empty[s], ph ^ 1, 21);
            if (elect_one()) {
                if SRB_DBG(dbg_skip & 1) { mbar_arrive(&full[s]); }
                else {
                    mbar_arrive_expect_tx(&full[s], bytes);
                    uint8_t* st = s_st + s * G::STAGE_STRIDE;
                    tma_load_4d(st, tmap_x, &full[s], (tx * G::TW - G::PAD) * 4, ty * G::TH, u.x_c8_off, n);
                    tma_load_4d(st + G::X_STRIDE, tmap_dy, &full[s], tx * G::TW * 4, u.dy_c8_off, ty * G::TH - G::PAD, n);
                }
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
    } else if (warp == 1) {
        const uint32_t idesc = umma_idesc_f16_mn(128, NN, BF16 ? 1 : 0);
        const uint64_t a_desc0 = umma_desc_nosw(0, 128, G::DYGRP);       // dY: groups (s, plane) 256 B apart
        const uint64_t b_desc0 = umma_desc_nosw(0, 128, G::XPLANE);      // X: channel planes
        int s = 0; uint32_t ph = 0; int it = 0;
        for (int tile = part; tile < u.num_tiles; tile += splitk, ++it) {
            mbar_wait(&full[s], ph, 22);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t xa = smem_u32(s_st + s * G::STAGE_STRIDE);
                const uint64_t a_desc = a_desc0 + (uint64_t)((xa + G::X_STRIDE) >> 4);
                const uint64_t b_desc = b_desc0 + (uint64_t)(xa >> 4);
                if (!SRB_DBG(dbg_skip & 2)) {
#pragma unroll 1
                    for (int r = 0; r < G::TH; ++r) {
                        const uint64_t ad = a_desc + (uint64_t)((r * G::DYROWP) >> 4);
#pragma unroll
                        for (int dx = 0; dx < KS; ++dx)
                            umma_f16_ss(tmem_base + dx * NN, ad, b_desc + (uint64_t)((r * G::XROWP + dx * 16) >> 4), idesc, (it | r) != 0);
                    }
                }
                umma_commit(&empty[s]);
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
        if (elect_one()) umma_commit(done);
        __syncwarp();
    } else if (warp < 4) {
        // ===== bias gradient db[co] = sum_p dY[p][co]: warps 2, 3 sum the (non-halo rows of the) dY tiles from shared memory
        if (KS == 3 && u.db != nullptr) {
            const int t = (warp - 2) * 32 + lane, plane = t >> 4, px = t & 15;
            float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            int s = 0; uint32_t ph = 0;
            for (int tile = part; tile < u.num_tiles; tile += splitk) {
                mbar_wait(&full[s], ph, 24);
                const uint32_t base = smem_u32(s_st + s * G::STAGE_STRIDE + G::X_STRIDE) + plane * G::DYGRP + px * 16;
#pragma unroll
                for (int r = 1; r <= G::TH; ++r) {
                    uint32_t q0, q1, q2, q3;
                    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3) : "r"(base + r * G::DYROWP));
                    const uint32_t qq[4] = {q0, q1, q2, q3};
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float2 f;
                        if constexpr (BF16) f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&qq[k]));
                        else f = __half22float2(*reinterpret_cast<const __half2*>(&qq[k]));
                        acc[2 * k] += f.x; acc[2 * k + 1] += f.y;
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty[s]);
                if (++s == G::STAGES) { s = 0; ph ^= 1; }
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float v = acc[k];
#pragma unroll
                for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (px == 0 && plane * 8 + k < u.cout_blk && part < u.num_tiles && !SRB_DBG(dbg_skip & 1)) atomicAdd(u.db + plane * 8 + k, v * u.scale);
            }
        }
    } else {
        // epilogue: once, after every tile of this CTA has been accumulated.  TMEM lane = s * NOUT + co (dy = KS-1 - s), column = ci
        const int lg = warp & 3;
        mbar_wait(done, 0, 23);
        tc_fence_after();
        const bool has_tiles = part < u.num_tiles && !SRB_DBG(dbg_skip & 2);
        const int L = lg * 32 + lane, sft = L / G::NOUT, co = L % G::NOUT;
        if (lg * 32 < KS * G::NOUT) {
            const bool lane_ok = has_tiles && sft < KS && co < u.cout_blk;
            float* dst0 = u.dw + ((size_t)(u.cout0 + co) * u.cin_total + u.cin0) * (KS * KS) + (KS - 1 - sft) * KS;
#pragma unroll 1
            for (int dx = 0; dx < KS; ++dx)
#pragma unroll 1
                for (int c0 = 0; c0 < u.cin_blk; c0 += 32) {
                    uint32_t rr[32];
                    tmem_ld32(tmem_base + dx * NN + c0 + ((uint32_t)(lg * 32) << 16), rr);
                    tmem_ld_wait();
                    if (lane_ok) {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (c0 + j < u.cin_blk) atomicAdd(dst0 + (size_t)(c0 + j) * (KS * KS) + dx, __uint_as_float(rr[j]) * u.scale);
                    }
                }
        }
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

// one conv, its template unit passed by value (srb_conv2d_wgrad, srb_conv9_wgrad): blockIdx.y enumerates the conv's
// (input-channel block, output-channel block) units, blockIdx.x the split-K parts of a unit's pixel tiles
template <bool BF16, int KS>
__global__ void __launch_bounds__(256, 1) wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmap_x,
                                                          const __grid_constant__ CUtensorMap tmap_dy, const WgradUnit u0, int cin,
                                                          int cout, int blk, int dbg_skip) {
    using G = WgradGeomT<KS>;
    const int ncb = (cout + G::NOUT - 1) / G::NOUT;
    const int cib = (int)blockIdx.y / ncb, cob = (int)blockIdx.y - cib * ncb;
    WgradUnit u = u0;
    u.cin0 = cib * blk; u.cout0 = cob * G::NOUT;
    u.cin_blk = cin - u.cin0 < blk ? cin - u.cin0 : blk;
    u.cout_blk = cout - u.cout0 < G::NOUT ? cout - u.cout0 : G::NOUT;
    u.x_c8_off += u.cin0 / 8; u.dy_c8_off += u.cout0 / 8;
    wgrad_body<BF16, KS>(&tmap_x, &tmap_dy, u, (int)blockIdx.x, (int)gridDim.x, dbg_skip);
}

// bias gradient: db[c] += scale * sum_p dY[p][c] over a C8 plane slice (HBM-bound, one pass, 16-byte loads)
template <bool BF16>
__global__ void __launch_bounds__(256) bias_grad_kernel(const uint4* __restrict__ dy, int c8_total, int c8_off, int planes,
                                                        size_t HW, int N, float scale, float* __restrict__ db, int cout) {
    // grid.y = plane, grid.x = chunks of pixels; every thread accumulates 8 channel sums
    const int pl = blockIdx.y;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const size_t total = (size_t)N * HW;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t n = i / HW, px = i - n * HW;
        const uint4 q = dy[(n * c8_total + c8_off + pl) * HW + px];
        const uint32_t qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float2 f;
            if constexpr (BF16) f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&qq[k]));
            else f = __half22float2(*reinterpret_cast<const __half2*>(&qq[k]));
            acc[2 * k] += f.x; acc[2 * k + 1] += f.y;
        }
    }
    __shared__ float s_acc[8][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) s_acc[threadIdx.x >> 5][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < 8) {
        float v = 0.f;
        for (int w = 0; w < 8; ++w) v += s_acc[w][threadIdx.x];
        const int c = pl * 8 + threadIdx.x;
        if (c < cout) atomicAdd(db + c, v * scale);
    }
}

// ------------------------------------------------------------------------------------------
// Batched weight gradients: every (conv, input-channel block, 32-channel cout block) of a whole backward pass is a
// UNIT of one launch.  The weight gradients are off the critical path of backpropagation (only the data gradients
// chain), so the training step defers them to the end and runs them as one grid of units x splitk CTAs (units sorted
// by cost, longest first, so the block scheduler packs the 148 SMs), instead of a 148-way split-K of one small filter
// at a time.  Tensor maps live in a device array (one pair per conv and channel-block width).
// ------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(256, 1) wgrad_batch_kernel(const CUtensorMap* __restrict__ tmaps, const WgradUnit* __restrict__ units,
                                                             int splitk, int dbg_skip) {
    const int uidx = blockIdx.x / splitk, part = blockIdx.x - uidx * splitk;
    const WgradUnit u = units[uidx];
    wgrad_body<BF16, 3>(tmaps + 2 * u.tm, tmaps + 2 * u.tm + 1, u, part, splitk, dbg_skip);
}

// batched bias gradients of the convs WITHOUT a 3x3 weight-gradient unit (9x9 head / tail): blockIdx.x = (conv, 8-channel
// plane), blockIdx.y = a slice of its pixels (one CTA per plane reads at ~15 GB/s: 0.3 ms for the 4 MB plane of the tail);
// partial sums are added with red.global.add (db is zeroed with the gradient arena)
struct BiasUnit {
    const uint4* dy; int c8_total, plane; int N; size_t HW;
    float scale; float* db; int nch;   // db points at the plane's first channel; nch real channels in it
};

template <bool BF16>
__global__ void __launch_bounds__(256) bias_grad_batch_kernel(const BiasUnit* __restrict__ units) {
    const BiasUnit u = units[blockIdx.x];
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int n = 0; n < u.N; ++n) {
        const uint4* base = u.dy + ((size_t)n * u.c8_total + u.plane) * u.HW;
        for (size_t px = blockIdx.y * 256u + threadIdx.x; px < u.HW; px += 256u * gridDim.y) {
            const uint4 q = base[px];
            const uint32_t qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float2 f;
                if constexpr (BF16) f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&qq[k]));
                else f = __half22float2(*reinterpret_cast<const __half2*>(&qq[k]));
                acc[2 * k] += f.x; acc[2 * k + 1] += f.y;
            }
        }
    }
    __shared__ float s_acc[8][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) s_acc[threadIdx.x >> 5][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < 8 && (int)threadIdx.x < u.nch) {
        float v = 0.f;
        for (int w = 0; w < 8; ++w) v += s_acc[w][threadIdx.x];
        atomicAdd(u.db + threadIdx.x, v * u.scale);
    }
}

}  // namespace srb
