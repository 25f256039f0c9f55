// HBM-bound helper kernels: input normalisation/packing, layout converters, L1 loss fwd/bwd.
// All are coalesced, 16-byte vectorised where the layout allows, grid sized in multiples of the SM count.
#pragma once
#include "srb_ptx.cuh"

namespace srb {

template <bool BF16>
__device__ __forceinline__ uint32_t pk2(float a, float b) {
    if constexpr (BF16) { __nv_bfloat162 h = __floats2bfloat162_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
    else { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
}
template <bool BF16>
__device__ __forceinline__ float ld16f(const void* p, size_t i) {
    if constexpr (BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
    else return __half2float(reinterpret_cast<const __half*>(p)[i]);
}

// Normalize.forward on uint8 (utils/datasets.py:65-71 as traced): f32(x)/255 - mean, /std — three
// separately rounded fp32 ops like the reference — then one rounding to the 16-bit operand type.
// in: [N][3][H][W] u8, out: [N][1][H][W][8] (channels 3..7 = 0).  One thread per pixel.
template <bool BF16>
__global__ void u8nchw_to_c8_kernel(const uint8_t* __restrict__ in, uint4* __restrict__ out, float m0, float m1,
                                    float m2, float s0, float s1, float s2, float maxv, size_t HW, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t n = i / HW, px = i - n * HW;
        const uint8_t* b = in + n * 3 * HW + px;
        float r = __fdiv_rn(__fsub_rn(__fdiv_rn((float)b[0], maxv), m0), s0);
        float g = __fdiv_rn(__fsub_rn(__fdiv_rn((float)b[HW], maxv), m1), s1);
        float bl = __fdiv_rn(__fsub_rn(__fdiv_rn((float)b[2 * HW], maxv), m2), s2);
        uint4 o;
        o.x = pk2<BF16>(r, g); o.y = pk2<BF16>(bl, 0.f); o.z = 0u; o.w = 0u;
        out[i] = o;
    }
}

// four pixels per thread (H*W % 4 == 0 and a 4-byte aligned input): one 32-bit load per colour plane, four 16-byte stores
template <bool BF16>
__global__ void u8nchw_to_c8_x4_kernel(const uint8_t* __restrict__ in, uint4* __restrict__ out, float m0, float m1,
                                       float m2, float s0, float s1, float s2, float maxv, unsigned HW4, size_t total4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total4; i += (size_t)gridDim.x * blockDim.x) {
        const size_t n = i / HW4;
        const unsigned q = (unsigned)(i - n * HW4);
        const uint32_t* b = reinterpret_cast<const uint32_t*>(in + n * 3 * (size_t)HW4 * 4) + q;
        const uint32_t r4 = __ldg(b), g4 = __ldg(b + HW4), b4 = __ldg(b + 2 * (size_t)HW4);
        uint4* o = out + i * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float r = __fdiv_rn(__fsub_rn(__fdiv_rn((float)((r4 >> (8 * k)) & 255u), maxv), m0), s0);
            const float g = __fdiv_rn(__fsub_rn(__fdiv_rn((float)((g4 >> (8 * k)) & 255u), maxv), m1), s1);
            const float bl = __fdiv_rn(__fsub_rn(__fdiv_rn((float)((b4 >> (8 * k)) & 255u), maxv), m2), s2);
            uint4 v;
            v.x = pk2<BF16>(r, g); v.y = pk2<BF16>(bl, 0.f); v.z = 0u; v.w = 0u;
            o[k] = v;
        }
    }
}

// Normalize into PIXEL-PAIR granules for the MODE_PAIR4 head conv: out[n][y][xs], xs = 0 .. W, = [pixel xs-1 | pixel xs]
// (4 x 16-bit each: c0 c1 c2 0; pixels -1 and W are zero).  One thread per granule.
template <bool BF16>
__global__ void u8nchw_to_pair4_kernel(const uint8_t* __restrict__ in, uint4* __restrict__ out, float m0, float m1,
                                       float m2, float s0, float s1, float s2, float maxv, int H, int W, size_t total) {
    const size_t HW = (size_t)H * W;
    const unsigned W1 = (unsigned)W + 1u;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t row = i / W1;                       // n * H + y
        const int xs = (int)(i - row * W1);
        const size_t n = row / (size_t)H, y = row - n * (size_t)H;
        const uint8_t* b = in + n * 3 * HW + y * (size_t)W;
        uint32_t h[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int x = xs - 1 + k;
            if (x >= 0 && x < W) {
                const float r = __fdiv_rn(__fsub_rn(__fdiv_rn((float)b[x], maxv), m0), s0);
                const float g = __fdiv_rn(__fsub_rn(__fdiv_rn((float)b[HW + x], maxv), m1), s1);
                const float bl = __fdiv_rn(__fsub_rn(__fdiv_rn((float)b[2 * HW + x], maxv), m2), s2);
                h[2 * k] = pk2<BF16>(r, g); h[2 * k + 1] = pk2<BF16>(bl, 0.f);
            }
        }
        out[i] = make_uint4(h[0], h[1], h[2], h[3]);
    }
}

// four pixels per thread (W % 4 == 0, 4-byte aligned input, blockDim a multiple of 32): every pixel's 8-byte half is
// normalised once; a granule needs the half of the pixel to its left, which comes from the neighbouring lane (lane 0 and
// the first thread of a row compute / zero it themselves).  Four 16-byte stores per thread, five at the end of a row.
template <bool BF16>
__global__ void u8nchw_to_pair4_x4_kernel(const uint8_t* __restrict__ in, uint4* __restrict__ out, float m0, float m1,
                                          float m2, float s0, float s1, float s2, float maxv, int H, int W, size_t total4) {
    const size_t HW = (size_t)H * W;
    const unsigned W4 = (unsigned)W / 4u;
    const int lane = threadIdx.x & 31;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i0 = blockIdx.x * (size_t)blockDim.x + (threadIdx.x & ~31u); i0 < total4; i0 += stride) {   // warp-uniform trip count
        const size_t i = i0 + lane;
        const bool act = i < total4;
        const size_t ii = act ? i : total4 - 1;
        const size_t row = ii / W4;                       // n * H + y
        const unsigned x0 = (unsigned)(ii - row * W4) * 4u;
        const size_t n = row / (size_t)H, y = row - n * (size_t)H;
        const uint8_t* b = in + n * 3 * HW + y * (size_t)W + x0;
        const uint32_t r4 = __ldg(reinterpret_cast<const uint32_t*>(b)), g4 = __ldg(reinterpret_cast<const uint32_t*>(b + HW)),
                       b4 = __ldg(reinterpret_cast<const uint32_t*>(b + 2 * HW));
        auto norm = [&](uint32_t r, uint32_t g, uint32_t bl) {
            const float fr = __fdiv_rn(__fsub_rn(__fdiv_rn((float)r, maxv), m0), s0);
            const float fg = __fdiv_rn(__fsub_rn(__fdiv_rn((float)g, maxv), m1), s1);
            const float fb = __fdiv_rn(__fsub_rn(__fdiv_rn((float)bl, maxv), m2), s2);
            return make_uint2(pk2<BF16>(fr, fg), pk2<BF16>(fb, 0.f));
        };
        uint2 h[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) h[k] = norm((r4 >> (8 * k)) & 255u, (g4 >> (8 * k)) & 255u, (b4 >> (8 * k)) & 255u);
        uint2 left;
        left.x = __shfl_up_sync(0xffffffffu, h[3].x, 1);
        left.y = __shfl_up_sync(0xffffffffu, h[3].y, 1);
        if (x0 == 0) left = make_uint2(0u, 0u);
        else if (lane == 0) left = norm(b[-1], b[HW - 1], b[2 * HW - 1]);
        if (!act) continue;
        uint4* o = out + row * ((size_t)W + 1) + x0;       // granule x0 = [pixel x0-1 | pixel x0]
        o[0] = make_uint4(left.x, left.y, h[0].x, h[0].y);
#pragma unroll
        for (int k = 1; k < 4; ++k) o[k] = make_uint4(h[k - 1].x, h[k - 1].y, h[k].x, h[k].y);
        if (x0 + 4 == (unsigned)W) o[4] = make_uint4(h[3].x, h[3].y, 0u, 0u);
    }
}

// fp32 NCHW [N][C][H][W] -> planes c8_off.. of a C8 buffer; channels beyond C in the last plane are zero.
template <bool BF16>
__global__ void f32nchw_to_c8_kernel(const float* __restrict__ in, uint4* __restrict__ out, int C, int c8_total,
                                     int c8_off, size_t HW, size_t total /* N * ceil(C/8) * HW */) {
    const int planes = (C + 7) / 8;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t px = i % HW; size_t t = i / HW; int pl = (int)(t % planes); size_t n = t / planes;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            int c = pl * 8 + j;
            v[j] = c < C ? in[(n * C + c) * HW + px] : 0.f;
        }
        uint4 o;
        o.x = pk2<BF16>(v[0], v[1]); o.y = pk2<BF16>(v[2], v[3]); o.z = pk2<BF16>(v[4], v[5]); o.w = pk2<BF16>(v[6], v[7]);
        out[(n * c8_total + c8_off + pl) * HW + px] = o;
    }
}

template <bool BF16>
__global__ void c8_to_f32nchw_kernel(const void* __restrict__ in, float* __restrict__ out, int C, int c8_total,
                                     int c8_off, size_t HW, size_t total /* N * C * HW */) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        size_t px = i % HW; size_t t = i / HW; int c = (int)(t % C); size_t n = t / C;
        out[i] = ld16f<BF16>(in, ((n * c8_total + c8_off + (c >> 3)) * HW + px) * 8 + (c & 7));
    }
}

// ------------------------------------------------------------------------------------------
// L1 loss (utils/loss.py:38-39).  Forward: per-thread fp32 partial sums over 16-byte loads,
// warp-shuffle tree, one partial per CTA, and the last CTA to finish adds the partials in index
// order (deterministic, no float atomics).  Backward: sign(p - t) * scale, sign(0) = 0.
// ------------------------------------------------------------------------------------------
constexpr int L1_MAX_BLOCKS = 1184;  // 148 SMs x 8
__device__ unsigned int g_l1_done_counter;

template <int DT>
__device__ __forceinline__ float ldp(const void* p, size_t i) {
    if constexpr (DT == 0) return __half2float(reinterpret_cast<const __half*>(p)[i]);
    else if constexpr (DT == 1) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
    else return reinterpret_cast<const float*>(p)[i];
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// SQ: squared error (nn.MSELoss, train.py:264 without --enchant) instead of the absolute error
template <bool SQ>
__device__ __forceinline__ float loss_term(float d) { return SQ ? d * d : fabsf(d); }

template <int DT, bool SQ = false>
__global__ void __launch_bounds__(256) l1_fwd_kernel(const void* __restrict__ pred, const float* __restrict__ tgt,
                                                     size_t n, float scale /* w / n */, float* __restrict__ loss,
                                                     float* __restrict__ partials) {
    float acc = 0.f;
    const size_t n4 = n / 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 t = reinterpret_cast<const float4*>(tgt)[i];
        float p0, p1, p2, p3;
        if constexpr (DT == 2) {
            float4 q = reinterpret_cast<const float4*>(pred)[i];
            p0 = q.x; p1 = q.y; p2 = q.z; p3 = q.w;
        } else {
            uint2 q = reinterpret_cast<const uint2*>(pred)[i];
            if constexpr (DT == 0) {
                float2 a = __half22float2(*reinterpret_cast<__half2*>(&q.x)), b = __half22float2(*reinterpret_cast<__half2*>(&q.y));
                p0 = a.x; p1 = a.y; p2 = b.x; p3 = b.y;
            } else {
                float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&q.x)), b = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&q.y));
                p0 = a.x; p1 = a.y; p2 = b.x; p3 = b.y;
            }
        }
        acc += loss_term<SQ>(p0 - t.x) + loss_term<SQ>(p1 - t.y) + loss_term<SQ>(p2 - t.z) + loss_term<SQ>(p3 - t.w);
    }
    for (size_t i = n4 * 4 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += stride)
        acc += loss_term<SQ>(ldp<DT>(pred, i) - tgt[i]);

    __shared__ float s_part[8];
    __shared__ bool s_last;
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < (blockDim.x >> 5) ? s_part[threadIdx.x] : 0.f;
        v = warp_sum(v);
        if (threadIdx.x == 0) {
            partials[blockIdx.x] = v;
            __threadfence();
            unsigned int done = atomicAdd(&g_l1_done_counter, 1u);
            s_last = (done == gridDim.x - 1);
        }
    }
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        __threadfence();
        double tot = 0.0;  // few hundred partials: serial-in-lane then fixed-order lane tree
        for (unsigned int b = threadIdx.x; b < gridDim.x; b += 32) tot += (double)reinterpret_cast<volatile float*>(partials)[b];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        if (threadIdx.x == 0) {
            *loss = (float)(tot * (double)scale);
            g_l1_done_counter = 0;
        }
    }
}

template <int DT, bool SQ = false>
__global__ void __launch_bounds__(256) l1_bwd_kernel(const void* __restrict__ pred, const float* __restrict__ tgt,
                                                     size_t n, float g /* w*scale/n */, void* __restrict__ dpred,
                                                     const float* __restrict__ g_dev /* optional device factor */) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    if (g_dev != nullptr) g *= *g_dev;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += stride) {
        float d = ldp<DT>(pred, i) - tgt[i];
        float s = SQ ? 2.f * d * g : (d > 0.f ? g : (d < 0.f ? -g : 0.f));
        if constexpr (DT == 0) reinterpret_cast<__half*>(dpred)[i] = __float2half_rn(s);
        else if constexpr (DT == 1) reinterpret_cast<__nv_bfloat16*>(dpred)[i] = __float2bfloat16_rn(s);
        else reinterpret_cast<float*>(dpred)[i] = s;
    }
}

}  // namespace srb
