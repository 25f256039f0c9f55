// Backward-pass helpers of the training step (train.py:57 autograd through utils/models.py:645-650):
// device-side operand packing from fp32 master weights, the elementwise backward glue that cannot be an
// epilogue (tanh', PixelShuffle inverse, gradient fan-in) and the 9x9 weight gradients of the head / tail convs.
#pragma once
#include "srb_ptx.cuh"
#include "srb_elem.cuh"

namespace srb {

// ------------------------------------------------------------------------------------------
// packed UMMA B operand from fp32 OIHW weights on the device (same element order as srb_pack_conv)
// ------------------------------------------------------------------------------------------
struct PackParams {
    int mode, c8in, nout, kblocks, cout, cin, k;
    int src_i, o_off, c_off, transpose;
    size_t total;   // 16-bit elements of the packed operand
};

template <bool BF16>
__device__ __forceinline__ void pack_conv_element(const float* __restrict__ w, uint16_t* __restrict__ pack, const PackParams& p,
                                                  const float* __restrict__ bias_src, float* __restrict__ bias_dst, size_t e) {
    if (e < (size_t)p.nout) bias_dst[e] = (bias_src != nullptr && (int)e < p.cout) ? bias_src[p.o_off + e] : 0.f;
    if (e >= p.total) return;
    const int step = 2 * p.nout * 8;
    const int ks = (int)(e / step);
    int rem = (int)(e - (size_t)ks * step);
    const int g = rem / (p.nout * 8);
    rem -= g * p.nout * 8;
    const int n = rem >> 3, j = rem & 7;
    int o = n, c = 0, dy = 0, dx = 0;
    bool valid = true;
    if (p.mode == 0) {   // MODE_STD
        if (p.kblocks) {   // [K block of 32 channels][dy][dx][2 K-steps]
            const int c2 = ks & 1; int t = ks >> 1;
            dx = t % 3; t /= 3; dy = t % 3; const int kb = t / 3;
            c = (kb * 4 + 2 * c2 + g) * 8 + j;
        } else {
            const int h = p.c8in / 2;
            const int c2 = ks % h; const int t = ks / h;
            dx = t % 3; dy = t / 3;
            c = (2 * c2 + g) * 8 + j;
        }
        valid = n < p.cout;
    } else if (p.mode == 1) {   // MODE_PAIRDX: the two K granules are neighbouring pixels
        dy = ks / 5; dx = 2 * (ks % 5) + g; c = j;
        valid = dx < 9 && j < p.cin && n < p.cout;
    } else {   // MODE_DXN: row n = dx*3 + co
        const int h = p.c8in / 2;
        dy = ks / h; const int c2 = ks % h;
        dx = n / 3; o = n - dx * 3;
        c = (2 * c2 + g) * 8 + j;
        valid = n < 27;
    }
    float v = 0.f;
    if (valid) {
        const int kk = p.k * p.k;
        if (p.transpose) v = w[((size_t)(p.c_off + c) * p.src_i + p.o_off + o) * kk + (p.k - 1 - dy) * p.k + (p.k - 1 - dx)];
        else v = w[((size_t)(p.o_off + o) * p.src_i + p.c_off + c) * kk + dy * p.k + dx];
    }
    uint16_t r;
    if constexpr (BF16) { __nv_bfloat16 h = __float2bfloat16_rn(v); r = *reinterpret_cast<uint16_t*>(&h); }
    else { __half h = __float2half_rn(v); r = *reinterpret_cast<uint16_t*>(&h); }
    pack[e] = r;
}

template <bool BF16>
__global__ void __launch_bounds__(256) pack_conv_kernel(const float* __restrict__ w, uint16_t* __restrict__ pack, PackParams p,
                                                        const float* __restrict__ bias_src, float* __restrict__ bias_dst) {
    pack_conv_element<BF16>(w, pack, p, bias_src, bias_dst, blockIdx.x * (size_t)blockDim.x + threadIdx.x);
}

// all operands of a net in ONE launch (the training step re-derives ~800 packed filters from the fp32 masters)
struct PackJob {
    PackParams p;
    const float* w; uint16_t* pack; const float* bias_src; float* bias_dst;
    int bf16;
};
__global__ void __launch_bounds__(256) pack_conv_batch_kernel(const PackJob* __restrict__ jobs) {
    const PackJob& j = jobs[blockIdx.y];
    for (size_t e = blockIdx.x * (size_t)blockDim.x + threadIdx.x; e < j.p.total; e += (size_t)gridDim.x * blockDim.x) {
        if (j.bf16) pack_conv_element<true>(j.w, j.pack, j.p, j.bias_src, j.bias_dst, e);
        else pack_conv_element<false>(j.w, j.pack, j.p, j.bias_src, j.bias_dst, e);
    }
}

// ------------------------------------------------------------------------------------------
// tanh backward of the tail conv (utils/models.py:636 act=Tanh): g_pre = g_out * (1 - out^2), fp32 NCHW [N,3,H,W]
// -> one C8 plane (channels 3..7 zero), the input of the tail conv's data/weight gradient.
// ------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void tanh_bwd_to_c8_kernel(const float* __restrict__ gout, const float* __restrict__ out, uint4* __restrict__ g8,
                                      float scale, size_t HW, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t n = i / HW, px = i - n * HW;
        const size_t b = n * 3 * HW + px;
        float v[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float o = out[b + c * HW];
            v[c] = gout[b + c * HW] * (1.f - o * o) * scale;
        }
        uint4 q;
        q.x = pk2<BF16>(v[0], v[1]); q.y = pk2<BF16>(v[2], 0.f); q.z = 0u; q.w = 0u;
        g8[i] = q;
    }
}

// ------------------------------------------------------------------------------------------
// Scaler backward glue (utils/models.py:580-583 reversed): u = LeakyReLU(PixelShuffle(2)(s)); given g_u
// [N][C/8][2H][2W][8] and the saved u, write g_s [N][4C/8][H][W][8]:
//   g_s[n, 4c + 2i + j, y, x] = g_u[n, c, 2y+i, 2x+j] * (u > 0 ? 1 : slope)
// One thread per (n, output plane k, y, x): conv channels 32k'.. — plane k of g_s holds channels 8k..8k+7 =
// (c = 2k, ij = 0..3), (c = 2k+1, ij = 0..3).
// ------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void pixunshuffle_mask_kernel(const void* __restrict__ gu, const void* __restrict__ u, uint4* __restrict__ gs,
                                         int C8 /* planes of g_u */, int H, int W, float slope, size_t total) {
    const size_t HW = (size_t)H * W, HW4 = HW * 4;
    const int W2 = 2 * W;
    const int planes_out = C8 * 4;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t px = i % HW; size_t t = i / HW;
        const int k = (int)(t % planes_out); const size_t n = t / planes_out;
        const int y = (int)(px / W), x = (int)(px - (size_t)y * W);
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int c = 2 * k + (q >> 2), ij = q & 3;
            const size_t src = ((n * C8 + (c >> 3)) * HW4 + (size_t)(2 * y + (ij >> 1)) * W2 + 2 * x + (ij & 1)) * 8 + (c & 7);
            const float g = ld16f<BF16>(gu, src), a = ld16f<BF16>(u, src);
            v[q] = a > 0.f ? g : g * slope;
        }
        uint4 o;
        o.x = pk2<BF16>(v[0], v[1]); o.y = pk2<BF16>(v[2], v[3]); o.z = pk2<BF16>(v[4], v[5]); o.w = pk2<BF16>(v[6], v[7]);
        gs[i] = o;
    }
}

// out = (alpha*a + beta*b) [* LeakyReLU'(act)] over C8 plane slices (gradient fan-in at the long skip, utils/models.py:647)
template <bool BF16>
__global__ void c8_axpby_mask_kernel(const uint4* __restrict__ a, int a_tot, int a_off, const uint4* __restrict__ b, int b_tot,
                                     int b_off, const uint4* __restrict__ act, int act_tot, int act_off, uint4* __restrict__ out,
                                     int o_tot, int o_off, int planes, float alpha, float beta, float slope, size_t HW, size_t total) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t px = i % HW; size_t t = i / HW;
        const int pl = (int)(t % planes); const size_t n = t / planes;
        const uint4 qa = a[(n * a_tot + a_off + pl) * HW + px];
        uint4 qb = make_uint4(0, 0, 0, 0), qm = make_uint4(0, 0, 0, 0);
        if (b != nullptr) qb = b[(n * b_tot + b_off + pl) * HW + px];
        if (act != nullptr) qm = act[(n * act_tot + act_off + pl) * HW + px];
        const uint32_t ua[4] = {qa.x, qa.y, qa.z, qa.w}, ub[4] = {qb.x, qb.y, qb.z, qb.w}, um[4] = {qm.x, qm.y, qm.z, qm.w};
        uint32_t r[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float2 fa, fb, fm;
            if constexpr (BF16) {
                fa = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ua[k]));
                fb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ub[k]));
                fm = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&um[k]));
            } else {
                fa = __half22float2(*reinterpret_cast<const __half2*>(&ua[k]));
                fb = __half22float2(*reinterpret_cast<const __half2*>(&ub[k]));
                fm = __half22float2(*reinterpret_cast<const __half2*>(&um[k]));
            }
            float v0 = alpha * fa.x + beta * fb.x, v1 = alpha * fa.y + beta * fb.y;
            if (act != nullptr) { if (!(fm.x > 0.f)) v0 *= slope; if (!(fm.y > 0.f)) v1 *= slope; }
            r[k] = pk2<BF16>(v0, v1);
        }
        out[(n * o_tot + o_off + pl) * HW + px] = make_uint4(r[0], r[1], r[2], r[3]);
    }
}

// ------------------------------------------------------------------------------------------
// Weight gradient of the 9x9 head (3->64) and tail (64->3) convs — 4 % of the FLOPs each, K = 3 is no tensor-core
// shape, so this one runs on the FP32 pipe:   T[c64][c3][tap] = sum_p A[p][c64] * S[p + sgn*(tap - centre)][c3]
//   head (utils/models.py:625): A = dY (64 ch), S = X (3 ch),  sgn = +1,  dW[c64][c3][tap]
//   tail (utils/models.py:636): A = X (64 ch),  S = dY (3 ch), sgn = -1,  dW[c3][c64][tap]
// A is a C8 8-plane slice, S a single C8 plane (channels 0..2).  Persistent CTAs walk 16x16 pixel tiles; thread
// (c64, tap group tg) keeps 21 taps x 3 channels in registers; the S reads are warp-uniform (broadcast).
// ------------------------------------------------------------------------------------------
struct Wgrad9Params {
    const void* a; int a_c8_total, a_c8_off;
    const void* s; int s_c8_total, s_c8_off;
    float* dw; float scale;
    int sgn, a_is_cout;
    int N, H, W, tiles_x, tiles_y, num_tiles;
};
constexpr int WG9_T = 16, WG9_SW = WG9_T + 8;
constexpr int WG9_SMEM = 256 * 64 * 4 + WG9_SW * WG9_SW * 16;

template <bool BF16>
__global__ void __launch_bounds__(256, 1) wgrad9_direct_kernel(const Wgrad9Params p) {
    extern __shared__ __align__(16) uint8_t wg9_smem[];
    float* s_a = reinterpret_cast<float*>(wg9_smem);                       // [256 px][64 ch]
    float4* s_s = reinterpret_cast<float4*>(wg9_smem + 256 * 64 * 4);      // [24][24] (3 ch + pad)
    const int tid = threadIdx.x;
    const int c64 = tid & 63, tg = tid >> 6;
    int off[21];
#pragma unroll
    for (int m = 0; m < 21; ++m) {
        const int tap = tg + 4 * m;
        const int dy = tap / 9, dx = tap - dy * 9;
        off[m] = tap < 81 ? p.sgn * ((dy - 4) * WG9_SW + (dx - 4)) : 0;
    }
    float acc[21][3];
#pragma unroll
    for (int m = 0; m < 21; ++m) { acc[m][0] = 0.f; acc[m][1] = 0.f; acc[m][2] = 0.f; }
    const size_t HW = (size_t)p.H * p.W;
    const int tiles_per_img = p.tiles_x * p.tiles_y;
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        const int n = tile / tiles_per_img;
        const int r = tile - n * tiles_per_img;
        const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
        __syncthreads();
        // A tile: 256 px x 8 planes, plane fastest across threads (conflict-free shared stores)
        for (int i = tid; i < 256 * 8; i += 256) {
            const int pl = i & 7, px = i >> 3;
            const int y = ty * WG9_T + (px >> 4), x = tx * WG9_T + (px & 15);
            float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (y < p.H && x < p.W) {
                const uint4 q = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(p.a) +
                    (((size_t)n * p.a_c8_total + p.a_c8_off + pl) * HW + (size_t)y * p.W + x) * 16);
                const uint32_t qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float2 f;
                    if constexpr (BF16) f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&qq[k]));
                    else f = __half22float2(*reinterpret_cast<const __half2*>(&qq[k]));
                    v[2 * k] = f.x; v[2 * k + 1] = f.y;
                }
            }
            float4* dst = reinterpret_cast<float4*>(s_a + px * 64 + pl * 8);
            dst[0] = make_float4(v[0], v[1], v[2], v[3]);
            dst[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
        for (int i = tid; i < WG9_SW * WG9_SW; i += 256) {
            const int sy = i / WG9_SW, sx = i - sy * WG9_SW;
            const int y = ty * WG9_T + sy - 4, x = tx * WG9_T + sx - 4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (y >= 0 && y < p.H && x >= 0 && x < p.W) {
                const uint2 q = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint8_t*>(p.s) +
                    (((size_t)n * p.s_c8_total + p.s_c8_off) * HW + (size_t)y * p.W + x) * 16);
                float2 f0, f1;
                if constexpr (BF16) {
                    f0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.x));
                    f1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&q.y));
                } else {
                    f0 = __half22float2(*reinterpret_cast<const __half2*>(&q.x));
                    f1 = __half22float2(*reinterpret_cast<const __half2*>(&q.y));
                }
                v = make_float4(f0.x, f0.y, f1.x, 0.f);
            }
            s_s[i] = v;
        }
        __syncthreads();
#pragma unroll 1
        for (int px = 0; px < 256; ++px) {
            const float a = s_a[px * 64 + c64];
            const int base = ((px >> 4) + 4) * WG9_SW + (px & 15) + 4;
#pragma unroll
            for (int m = 0; m < 21; ++m) {
                const float4 sv = s_s[base + off[m]];
                acc[m][0] = fmaf(a, sv.x, acc[m][0]);
                acc[m][1] = fmaf(a, sv.y, acc[m][1]);
                acc[m][2] = fmaf(a, sv.z, acc[m][2]);
            }
        }
    }
    if (blockIdx.x < p.num_tiles) {
#pragma unroll
        for (int m = 0; m < 21; ++m) {
            const int tap = tg + 4 * m;
            if (tap < 81) {
#pragma unroll
                for (int c3 = 0; c3 < 3; ++c3) {
                    const size_t idx = p.a_is_cout ? ((size_t)c64 * 3 + c3) * 81 + tap : ((size_t)c3 * 64 + c64) * 81 + tap;
                    atomicAdd(p.dw + idx, acc[m][c3] * p.scale);
                }
            }
        }
    }
}

}  // namespace srb
