// Train-mode BatchNorm2d of the reference's `Conv` block (utils/models.py:75-111: conv -> nn.BatchNorm2d -> act) over C8-plane slices:
// batch statistics, normalise + activation (+ the residual terms of the block that follows), and the backward reductions.
//
// A BatchNorm in training mode needs every pixel of the batch before any output (forward: mean / variance; backward: the two
// sums of dz and dz * x_hat), so it cannot live in the epilogue of the conv that produces its input: the conv stores its raw
// output y (16-bit), `bn_stats_kernel` reduces it, `bn_apply_kernel` normalises.  All four kernels are HBM-bound passes over a
// slice of P planes (P = 4 or 8 here: 32 or 64 channels) — algorithmic bytes per element: stats 2, apply 2 + 2 (+ 2 per residual),
// backward reduce 2 + 2, backward apply 2 + 2 + 2.
//
// Indexing: one block = 256 consecutive pixels of one (image, plane) per iteration, a thread = one 16-byte granule (8 channels of
// its plane), so all threads of a block work on the same 8 channels and the block reduces them with shuffles; one double
// atomicAdd per channel, sum and block.  (fp32 partial sums inside a thread / block, double across blocks: E[y^2] - mean^2 with
// a million elements per channel.)
#pragma once
#include <cstdint>

#include "srb_elem.cuh"

namespace srb {

struct BNParams {
    const void* y; int y_c8_total, y_c8_off;          // raw conv output (pre-BN)
    void* out; int out_c8_total, out_c8_off;          // destination of apply / bwd_apply
    int planes;                                       // channels / 8 of the slice
    const float* gamma; const float* beta;            // [C]
    double* sums;                                     // [2][C]: sum y, sum y^2 of this step (zeroed by the caller)
    float* running_mean; float* running_var; long long* nbt;
    float momentum, eps;
    float slope; int act;                             // act: 0 identity, 1 LeakyReLU(slope)
    float out_scale;
    const void* res; int res_c8_total, res_c8_off; float res_scale;
    const void* res2; int res2_c8_total, res2_c8_off; float res2_scale;
    const void* da; int da_c8_total, da_c8_off; float up_scale;      // backward: gradient of the block's output (after act)
    double* bsums;                                    // [2][C]: sum dz, sum dz * x_hat (zeroed by the caller)
    float* dgamma; float* dbeta; float grad_scale;    // fp32 parameter gradients, += grad_scale * sum
    int N, H, W;
};

template <bool BF16>
__device__ __forceinline__ void unpack8(const uint4& q, float* v) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if constexpr (BF16) {
            float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[k]));
            v[2 * k] = f.x; v[2 * k + 1] = f.y;
        } else {
            float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[k]));
            v[2 * k] = f.x; v[2 * k + 1] = f.y;
        }
    }
}
template <bool BF16>
__device__ __forceinline__ uint4 pack8(const float* v) {
    uint4 o;
    o.x = pk2<BF16>(v[0], v[1]); o.y = pk2<BF16>(v[2], v[3]); o.z = pk2<BF16>(v[4], v[5]); o.w = pk2<BF16>(v[6], v[7]);
    return o;
}

// block-wide sum of 8 + 8 per-thread values (all threads: same 8 channels); result valid in thread 0
__device__ __forceinline__ void block_sum16(float* a, float* b, float (*s_red)[16]) {
#pragma unroll
    for (int k = 0; k < 8; ++k) { a[k] = warp_sum(a[k]); b[k] = warp_sum(b[k]); }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) { s_red[warp][k] = a[k]; s_red[warp][8 + k] = b[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 16) {
        float t = 0.f;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_red[w][threadIdx.x];
        s_red[0][threadIdx.x] = t;
    }
    __syncthreads();
}

// grid: (chunks, planes, N); a block walks `HW / chunks` pixels of its (image, plane)
template <bool BF16>
__global__ void __launch_bounds__(256) bn_stats_kernel(const BNParams p) {
    __shared__ float s_red[8][16];
    const size_t HW = (size_t)p.H * p.W;
    const int pl = blockIdx.y, n = blockIdx.z;
    const uint4* y = reinterpret_cast<const uint4*>(p.y) + ((size_t)n * p.y_c8_total + p.y_c8_off + pl) * HW;
    float s[8] = {}, q[8] = {};
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < HW; i += (size_t)gridDim.x * blockDim.x) {
        float v[8];
        unpack8<BF16>(y[i], v);
#pragma unroll
        for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] += v[k] * v[k]; }
    }
    block_sum16(s, q, s_red);
    const int C = p.planes * 8;
    if (threadIdx.x < 8) atomicAdd(&p.sums[pl * 8 + threadIdx.x], (double)s_red[0][threadIdx.x]);
    else if (threadIdx.x < 16) atomicAdd(&p.sums[C + pl * 8 + threadIdx.x - 8], (double)s_red[0][threadIdx.x]);
}

struct BNCh { float mean, istd, g, b; };
__device__ __forceinline__ BNCh bn_channel(const BNParams& p, int c, double M) {
    const int C = p.planes * 8;
    const double mean = p.sums[c] / M;
    double var = p.sums[C + c] / M - mean * mean;
    if (var < 0.0) var = 0.0;
    BNCh r;
    r.mean = (float)mean; r.istd = (float)(1.0 / sqrt(var + (double)p.eps));
    r.g = p.gamma[c]; r.b = p.beta[c];
    return r;
}

// out = out_scale * act(gamma * x_hat + beta) + res_scale * res + res2_scale * res2;  block (0,0,0) also moves the running statistics
template <bool BF16>
__global__ void __launch_bounds__(256) bn_apply_kernel(const BNParams p) {
    const size_t HW = (size_t)p.H * p.W;
    const double M = (double)p.N * (double)HW;
    const int pl = blockIdx.y, n = blockIdx.z;
    __shared__ BNCh s_ch[8];
    if (threadIdx.x < 8) s_ch[threadIdx.x] = bn_channel(p, pl * 8 + threadIdx.x, M);
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && p.running_mean != nullptr) {
        // nn.BatchNorm2d in training mode: running = (1 - momentum) * running + momentum * batch (the variance unbiased)
        const int C = p.planes * 8;
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            const double mean = p.sums[c] / M;
            double var = p.sums[C + c] / M - mean * mean;
            if (var < 0.0) var = 0.0;
            const double unb = M > 1.0 ? var * M / (M - 1.0) : var;
            p.running_mean[c] = (1.f - p.momentum) * p.running_mean[c] + p.momentum * (float)mean;
            p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * (float)unb;
        }
        if (threadIdx.x == 0 && p.nbt != nullptr) *p.nbt += 1;
    }
    __syncthreads();
    const uint4* y = reinterpret_cast<const uint4*>(p.y) + ((size_t)n * p.y_c8_total + p.y_c8_off + pl) * HW;
    uint4* o = reinterpret_cast<uint4*>(p.out) + ((size_t)n * p.out_c8_total + p.out_c8_off + pl) * HW;
    const uint4* r1 = p.res ? reinterpret_cast<const uint4*>(p.res) + ((size_t)n * p.res_c8_total + p.res_c8_off + pl) * HW : nullptr;
    const uint4* r2 = p.res2 ? reinterpret_cast<const uint4*>(p.res2) + ((size_t)n * p.res2_c8_total + p.res2_c8_off + pl) * HW : nullptr;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < HW; i += (size_t)gridDim.x * blockDim.x) {
        float v[8];
        unpack8<BF16>(y[i], v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const BNCh c = s_ch[k];
            float z = (v[k] - c.mean) * c.istd * c.g + c.b;
            if (p.act == 1) z = z > 0.f ? z : z * p.slope;
            v[k] = z * p.out_scale;
        }
        if (r1) {
            float a[8];
            unpack8<BF16>(r1[i], a);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += p.res_scale * a[k];
        }
        if (r2) {
            float a[8];
            unpack8<BF16>(r2[i], a);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] += p.res2_scale * a[k];
        }
        o[i] = pack8<BF16>(v);
    }
}

// dz = up_scale * da * act'(z);  sums of dz and dz * x_hat per channel
template <bool BF16>
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const BNParams p) {
    __shared__ float s_red[8][16];
    __shared__ BNCh s_ch[8];
    const size_t HW = (size_t)p.H * p.W;
    const double M = (double)p.N * (double)HW;
    const int pl = blockIdx.y, n = blockIdx.z;
    if (threadIdx.x < 8) s_ch[threadIdx.x] = bn_channel(p, pl * 8 + threadIdx.x, M);
    __syncthreads();
    const uint4* y = reinterpret_cast<const uint4*>(p.y) + ((size_t)n * p.y_c8_total + p.y_c8_off + pl) * HW;
    const uint4* da = reinterpret_cast<const uint4*>(p.da) + ((size_t)n * p.da_c8_total + p.da_c8_off + pl) * HW;
    float s[8] = {}, q[8] = {};
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < HW; i += (size_t)gridDim.x * blockDim.x) {
        float v[8], g[8];
        unpack8<BF16>(y[i], v);
        unpack8<BF16>(da[i], g);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const BNCh c = s_ch[k];
            const float xh = (v[k] - c.mean) * c.istd;
            float dz = g[k] * p.up_scale;
            if (p.act == 1 && xh * c.g + c.b <= 0.f) dz *= p.slope;
            s[k] += dz; q[k] += dz * xh;
        }
    }
    block_sum16(s, q, s_red);
    const int C = p.planes * 8;
    if (threadIdx.x < 8) atomicAdd(&p.bsums[pl * 8 + threadIdx.x], (double)s_red[0][threadIdx.x]);
    else if (threadIdx.x < 16) atomicAdd(&p.bsums[C + pl * 8 + threadIdx.x - 8], (double)s_red[0][threadIdx.x]);
}

// dy = gamma * istd * (dz - mean(dz) - x_hat * mean(dz * x_hat)) -> out;  block (0,0,0): dgamma += grad_scale * sum(dz * x_hat), dbeta likewise
template <bool BF16>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const BNParams p) {
    __shared__ BNCh s_ch[8];
    __shared__ float s_m1[8], s_m2[8];
    const size_t HW = (size_t)p.H * p.W;
    const double M = (double)p.N * (double)HW;
    const int pl = blockIdx.y, n = blockIdx.z, C = p.planes * 8;
    if (threadIdx.x < 8) {
        s_ch[threadIdx.x] = bn_channel(p, pl * 8 + threadIdx.x, M);
        s_m1[threadIdx.x] = (float)(p.bsums[pl * 8 + threadIdx.x] / M);
        s_m2[threadIdx.x] = (float)(p.bsums[C + pl * 8 + threadIdx.x] / M);
    }
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            if (p.dbeta) p.dbeta[c] += p.grad_scale * (float)p.bsums[c];
            if (p.dgamma) p.dgamma[c] += p.grad_scale * (float)p.bsums[C + c];
        }
    }
    __syncthreads();
    const uint4* y = reinterpret_cast<const uint4*>(p.y) + ((size_t)n * p.y_c8_total + p.y_c8_off + pl) * HW;
    const uint4* da = reinterpret_cast<const uint4*>(p.da) + ((size_t)n * p.da_c8_total + p.da_c8_off + pl) * HW;
    uint4* o = reinterpret_cast<uint4*>(p.out) + ((size_t)n * p.out_c8_total + p.out_c8_off + pl) * HW;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < HW; i += (size_t)gridDim.x * blockDim.x) {
        float v[8], g[8];
        unpack8<BF16>(y[i], v);
        unpack8<BF16>(da[i], g);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const BNCh c = s_ch[k];
            const float xh = (v[k] - c.mean) * c.istd;
            float dz = g[k] * p.up_scale;
            if (p.act == 1 && xh * c.g + c.b <= 0.f) dz *= p.slope;
            v[k] = c.g * c.istd * (dz - s_m1[k] - xh * s_m2[k]);
        }
        o[i] = pack8<BF16>(v);
    }
}

}  // namespace srb
