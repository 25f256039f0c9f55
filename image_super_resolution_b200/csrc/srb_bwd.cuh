// Backward-pass helpers of the training step (train.py:57 autograd through utils/models.py:645-650):
// device-side operand packing from fp32 master weights (also operands gathered from several source filters: the
// data-gradient convs of a dense block) and the elementwise backward glue that cannot be an epilogue (tanh',
// PixelShuffle inverse, gradient fan-in).  Weight / bias gradients: srb_wgrad.cuh.
#pragma once
#include "srb_ptx.cuh"
#include "srb_elem.cuh"

namespace srb {

// ------------------------------------------------------------------------------------------
// packed UMMA B operand from fp32 OIHW weights on the device (same element order as srb_pack_conv)
// ------------------------------------------------------------------------------------------
struct PackParams {
    int mode, c8in, nout, nmma, kblocks, cout, cin, k;   // nmma: B rows per K granule (3 * nout for the dy-folded operands)
    int src_i, o_off, c_off, transpose;
    int c_lo, c_n;  // c_n > 0: this job fills only the packed input channels [c_lo, c_lo + c_n) (source channel = c - c_lo + c_off):
                    // an operand whose input channels are gathered from several source filters (one job per source)
    float scale;    // multiplies the source weights (a constant factor of the layer folded into the operand)
    size_t total;   // 16-bit elements of the packed operand
};

template <bool BF16>
__device__ __forceinline__ void pack_conv_element(const float* __restrict__ w, uint16_t* __restrict__ pack, const PackParams& p,
                                                  const float* __restrict__ bias_src, float* __restrict__ bias_dst, size_t e) {
    if (e < (size_t)p.nout) bias_dst[e] = (bias_src != nullptr && (int)e < p.cout) ? bias_src[p.o_off + e] : 0.f;
    if (e >= p.total) return;
    const int step = 2 * p.nmma * 8;
    const int ks = (int)(e / step);
    int rem = (int)(e - (size_t)ks * step);
    const int g = rem / (p.nmma * 8);
    rem -= g * p.nmma * 8;
    const int n = rem >> 3, j = rem & 7;
    int o = n, c = 0, dy = 0, dx = 0;
    bool valid = true;
    if (p.mode == 4) {   // MODE_DYF (weight streaming): K-steps [K block of 32 channels][dx][channel pair], B row = dy * nout + co
        const int c2 = ks & 1; const int t = ks >> 1;
        dx = t % 3; const int kb = t / 3;
        dy = n / p.nout; o = n - dy * p.nout;
        c = (kb * 4 + 2 * c2 + g) * 8 + j;
        valid = o < p.cout;
    } else if (p.mode == 0) {   // MODE_STD
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
    if (p.c_n > 0) {
        if (c < p.c_lo || c >= p.c_lo + p.c_n) return;     // another job's channels
        c -= p.c_lo;
    }
    float v = 0.f;
    if (valid) {
        const int kk = p.k * p.k;
        if (p.transpose) v = w[((size_t)(p.c_off + c) * p.src_i + p.o_off + o) * kk + (p.k - 1 - dy) * p.k + (p.k - 1 - dx)];
        else v = w[((size_t)(p.o_off + o) * p.src_i + p.c_off + c) * kk + dy * p.k + dx];
        v *= p.scale;
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

// bare nn.PixelShuffle(2) + LeakyReLU of the current Denoise generator (utils/models.py:685: residual_conv1), not preceded by a conv
// whose store could carry it (the conv before it adds a residual first): u[n, c, 2y+i, 2x+j] = lrelu(s[n, 4c+2i+j, y, x]).
// s: [N][4P][H][W][8], u: [N][P][2H][2W][8]; one thread per output granule.
template <bool BF16>
__global__ void pixshuffle_lrelu_kernel(const void* __restrict__ sbuf, uint4* __restrict__ u, int P, int H, int W, float slope, size_t total) {
    const size_t HW = (size_t)H * W;
    const int W2 = 2 * W, H2 = 2 * H;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int X = (int)(i % W2); size_t t = i / W2;
        const int Y = (int)(t % H2); t /= H2;
        const int k = (int)(t % P); const size_t n = t / P;
        const int ij = 2 * (Y & 1) + (X & 1);
        const size_t px = (size_t)(Y >> 1) * W + (X >> 1);
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int sc = 32 * k + 4 * q + ij;                        // source channel
            const float a = ld16f<BF16>(sbuf, ((n * (4 * (size_t)P) + (sc >> 3)) * HW + px) * 8 + (sc & 7));
            v[q] = a > 0.f ? a : a * slope;
        }
        uint4 o;
        o.x = pk2<BF16>(v[0], v[1]); o.y = pk2<BF16>(v[2], v[3]); o.z = pk2<BF16>(v[4], v[5]); o.w = pk2<BF16>(v[6], v[7]);
        u[i] = o;
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

// (the 9x9 head / tail weight gradients run on the folded tensor-core GEMM of srb_wgrad.cuh, KS = 9)

}  // namespace srb
