// Weight gradient of a 3x3 conv on tcgen05/TMEM:  dW[co][ci][dy][dx] = sum_p dY[p][co] * X[p + (dy-1, dx-1)][ci]
// (autograd of aten::_convolution reached from train.py:57).
//
// GEMM view per tap: D[ci][co] += A[ci][p] * B[co][p], reduction K = pixels.  Both operands are the SAME C8-plane
// tiles the forward kernel loads, now read "MN-major": a 16-byte granule holds 8 channels (contiguous M / N) of one
// pixel and consecutive pixels are 16 B apart (K), so 8 pixels x 8 channels is one 128-byte core matrix;
// LBO = 128 B (next 8 pixels of the row), SBO = plane pitch (next 8 channels).  A tap is a start-address shift of
// the X halo tile.  M = 128 input channels (16 planes; planes past the tensor are TMA zero-filled),
// N = 32 output channels, K = 16 pixels (one tile row) per MMA; the 9 taps own 9 TMEM accumulators that persist
// over all pixel tiles of the CTA, so consecutive MMAs are independent.  The CTA's partial dW is added to global
// fp32 with red.global.add (order not fixed: documented tolerance in the tests).
#pragma once
#include <cuda.h>
#include "srb_ptx.cuh"

namespace srb {

struct WgradParams {
    int N, H, W;
    int tiles_x, tiles_y, num_tiles;
    int x_c8_off, dy_c8_off;      // first plane of the cin block / cout block
    int cin_blk, cout_blk;        // real channels in this block (<= 128 / <= 32)
    int cin_total, cin0, cout0;   // dW indexing: [cout][cin_total][3][3]
    float* dw;
    float scale;
    int dbg_swap;
};

struct WgradGeom {
    static constexpr int TW = 16, TH = 8;
    static constexpr int XW = TW + 2, XH = TH + 2, XPLANES = 16, DYPLANES = 4;
    static constexpr int XROWP = XW * 16, XPLANE = XH * XROWP, XBYTES = XPLANES * XPLANE;       // 46080
    static constexpr int DYROWP = TW * 16, DYPLANE = TH * DYROWP, DYBYTES = DYPLANES * DYPLANE;  // 8192
    static constexpr int X_STRIDE = (XBYTES + 1023) / 1024 * 1024;
    static constexpr int STAGE_STRIDE = X_STRIDE + DYBYTES;
    static constexpr int STAGES = 4;
    static constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_STRIDE + 256;
    static constexpr int NOUT = 32;
};

// kind::f16 instruction descriptor with both operands MN-major
__host__ __device__ constexpr uint32_t umma_idesc_f16_mn(int M, int N, int fmt) {
    return umma_idesc_f16(M, N, fmt) | (1u << 15) | (1u << 16);
}

template <bool BF16>
__global__ void __launch_bounds__(256, 1) wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmap_x,
                                                          const __grid_constant__ CUtensorMap tmap_dy, const WgradParams p) {
    using G = WgradGeom;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* s_st = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::STAGES * G::STAGE_STRIDE);
    uint64_t* full = bars;          // [4]
    uint64_t* empty = bars + 4;     // [4]
    uint64_t* done = bars + 8;      // [1]
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 10);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < G::STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(done, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tmap_x);
        tma_prefetch_desc(&tmap_dy);
    }
    if (warp == 0) { __syncwarp(); tmem_alloc(s_tmem, 512); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    const int tiles_per_img = p.tiles_x * p.tiles_y;

    if (warp == 0) {
        int s = 0; uint32_t ph = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
            const int n = tile / tiles_per_img;
            const int r = tile - n * tiles_per_img;
            const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
            mbar_wait(&empty[s], ph ^ 1, 11);
            if (elect_one()) {
                mbar_arrive_expect_tx(&full[s], G::XBYTES + G::DYBYTES);
                uint8_t* st = s_st + s * G::STAGE_STRIDE;
                tma_load_4d(st, &tmap_x, &full[s], (tx * G::TW - 1) * 4, ty * G::TH - 1, p.x_c8_off, n);
                tma_load_4d(st + G::X_STRIDE, &tmap_dy, &full[s], tx * G::TW * 4, ty * G::TH, p.dy_c8_off, n);
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc = umma_idesc_f16_mn(128, G::NOUT, BF16 ? 1 : 0);
        uint32_t a_lbo = 128, a_sbo = G::XPLANE, b_lbo = 128, b_sbo = G::DYPLANE;
        if (p.dbg_swap) { uint32_t t = a_lbo; a_lbo = a_sbo; a_sbo = t; t = b_lbo; b_lbo = b_sbo; b_sbo = t; }
        const uint64_t a_desc0 = umma_desc_nosw(0, a_lbo, a_sbo);
        const uint64_t b_desc0 = umma_desc_nosw(0, b_lbo, b_sbo);
        int s = 0; uint32_t ph = 0; int it = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            mbar_wait(&full[s], ph, 12);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t xa = smem_u32(s_st + s * G::STAGE_STRIDE);
                const uint64_t a_desc = a_desc0 + (uint64_t)(xa >> 4);
                const uint64_t b_desc = b_desc0 + (uint64_t)((xa + G::X_STRIDE) >> 4);
#pragma unroll 1
                for (int r = 0; r < G::TH; ++r) {
                    const uint64_t bd = b_desc + (uint64_t)((r * G::DYROWP) >> 4);
#pragma unroll
                    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx)
                            umma_f16_ss(tmem_base + (dy * 3 + dx) * G::NOUT,
                                        a_desc + (uint64_t)(((r + dy) * G::XROWP + dx * 16) >> 4), bd, idesc, (it | r) != 0);
                }
                umma_commit(&empty[s]);
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
        if (elect_one()) umma_commit(done);
        __syncwarp();
    } else if (warp >= 4) {
        // epilogue: once, after every tile of this CTA has been accumulated
        const int lg = warp & 3;
        const int ci = lg * 32 + lane;
        mbar_wait(done, 0, 13);
        tc_fence_after();
        const bool has_tiles = blockIdx.x < p.num_tiles;
#pragma unroll 1
        for (int tap = 0; tap < 9; ++tap) {
            uint32_t rr[32];
            tmem_ld32(tmem_base + tap * G::NOUT + ((uint32_t)(lg * 32) << 16), rr);
            tmem_ld_wait();
            if (has_tiles && ci < p.cin_blk) {
                float* dst = p.dw + ((size_t)p.cout0 * p.cin_total + p.cin0 + ci) * 9 + tap;
#pragma unroll
                for (int co = 0; co < 32; ++co)
                    if (co < p.cout_blk) atomicAdd(dst + (size_t)co * p.cin_total * 9, __uint_as_float(rr[co]) * p.scale);
            }
        }
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
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
// Batched weight gradients: every (conv, 128-channel cin block, 32-channel cout block) of a whole backward pass is a
// UNIT of one launch.  The weight gradients are off the critical path of backpropagation (only the data gradients
// chain), so the training step defers them to the end and runs them as one grid of units x splitk CTAs: with
// splitk = 1 each dW element is produced by exactly one CTA (no cross-CTA reduction, no contended atomics) and the
// 148 SMs are kept busy by ~450 independent units instead of 148-way split-K of one small filter at a time.
// Tensor maps live in a device array (one pair per conv).
// ------------------------------------------------------------------------------------------
struct WgradUnit {
    int tm;                       // index of the conv's tensor-map pair: tmaps[2*tm] = X, tmaps[2*tm+1] = dY
    int x_c8_off, dy_c8_off;
    int cin_blk, cout_blk, cin_total, cin0, cout0;
    int N, H, W, tiles_x, tiles_y, num_tiles;
    float scale;
    float* dw;
};

template <bool BF16>
__global__ void __launch_bounds__(256, 1) wgrad_batch_kernel(const CUtensorMap* __restrict__ tmaps, const WgradUnit* __restrict__ units,
                                                             int splitk) {
    using G = WgradGeom;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* s_st = smem;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::STAGES * G::STAGE_STRIDE);
    uint64_t* full = bars;
    uint64_t* empty = bars + 4;
    uint64_t* done = bars + 8;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 10);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int uidx = blockIdx.x / splitk, part = blockIdx.x - uidx * splitk;
    const WgradUnit u = units[uidx];
    const CUtensorMap* tmap_x = tmaps + 2 * u.tm;
    const CUtensorMap* tmap_dy = tmap_x + 1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < G::STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
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

    if (warp == 0) {
        int s = 0; uint32_t ph = 0;
        for (int tile = part; tile < u.num_tiles; tile += splitk) {
            const int n = tile / tiles_per_img;
            const int r = tile - n * tiles_per_img;
            const int ty = r / u.tiles_x, tx = r - ty * u.tiles_x;
            mbar_wait(&empty[s], ph ^ 1, 21);
            if (elect_one()) {
                mbar_arrive_expect_tx(&full[s], G::XBYTES + G::DYBYTES);
                uint8_t* st = s_st + s * G::STAGE_STRIDE;
                tma_load_4d(st, tmap_x, &full[s], (tx * G::TW - 1) * 4, ty * G::TH - 1, u.x_c8_off, n);
                tma_load_4d(st + G::X_STRIDE, tmap_dy, &full[s], tx * G::TW * 4, ty * G::TH, u.dy_c8_off, n);
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc = umma_idesc_f16_mn(128, G::NOUT, BF16 ? 1 : 0);
        const uint64_t a_desc0 = umma_desc_nosw(0, 128, G::XPLANE);
        const uint64_t b_desc0 = umma_desc_nosw(0, 128, G::DYPLANE);
        int s = 0; uint32_t ph = 0; int it = 0;
        for (int tile = part; tile < u.num_tiles; tile += splitk, ++it) {
            mbar_wait(&full[s], ph, 22);
            tc_fence_after();
            if (elect_one()) {
                const uint32_t xa = smem_u32(s_st + s * G::STAGE_STRIDE);
                const uint64_t a_desc = a_desc0 + (uint64_t)(xa >> 4);
                const uint64_t b_desc = b_desc0 + (uint64_t)((xa + G::X_STRIDE) >> 4);
#pragma unroll 1
                for (int r = 0; r < G::TH; ++r) {
                    const uint64_t bd = b_desc + (uint64_t)((r * G::DYROWP) >> 4);
#pragma unroll
                    for (int dy = 0; dy < 3; ++dy)
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx)
                            umma_f16_ss(tmem_base + (dy * 3 + dx) * G::NOUT,
                                        a_desc + (uint64_t)(((r + dy) * G::XROWP + dx * 16) >> 4), bd, idesc, (it | r) != 0);
                }
                umma_commit(&empty[s]);
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
        if (elect_one()) umma_commit(done);
        __syncwarp();
    } else if (warp >= 4) {
        const int lg = warp & 3;
        const int ci = lg * 32 + lane;
        mbar_wait(done, 0, 23);
        tc_fence_after();
        const bool has_tiles = part < u.num_tiles;
#pragma unroll 1
        for (int tap = 0; tap < 9; ++tap) {
            uint32_t rr[32];
            tmem_ld32(tmem_base + tap * G::NOUT + ((uint32_t)(lg * 32) << 16), rr);
            tmem_ld_wait();
            if (has_tiles && ci < u.cin_blk) {
                float* dst = u.dw + ((size_t)u.cout0 * u.cin_total + u.cin0 + ci) * 9 + tap;
#pragma unroll
                for (int co = 0; co < 32; ++co)
                    if (co < u.cout_blk) atomicAdd(dst + (size_t)co * u.cin_total * 9, __uint_as_float(rr[co]) * u.scale);
            }
        }
        tc_fence_before();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

// batched bias gradients: one CTA per (conv, 8-channel plane); fixed-order in-CTA reduction, plain store (deterministic)
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
        for (size_t px = threadIdx.x; px < u.HW; px += 256) {
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
        u.db[threadIdx.x] += v * u.scale;
    }
}

}  // namespace srb
