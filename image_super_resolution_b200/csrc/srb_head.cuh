// Head of the model.pt graph in ONE launch: Normalize (utils/datasets.py:65-71) + conv0 9x9 3 -> 64 + bias + LeakyReLU
// (utils/models.py:97-98, fold :366-406) straight from the uint8 frame — SURVEY §8 a1 + a2 and the ingest half of §8(f) rank 2
// (uint8 NCHW / HWC frame ingest folded into conv0's loader).  No normalised input plane ever exists in HBM.
//
// The A operand is BUILT in shared memory by producer warps (generic proxy, then fence.proxy.async) instead of being
// TMA-loaded, which frees its layout from anything a tensor map can express:
//   * a K = 16 instruction takes FIVE neighbouring pixels x 3 channels (+ 1 zero): K-half 0 is a granule of plane A
//     [p0.c0 c1 c2 | p1.c0 c1 c2 | p2.c0 c1], K-half 1 the granule of plane B at the same position [p2.c2 | p3.c0 c1 c2 |
//     p4.c0 c1 c2 | 0].  Position g of a halo row holds the granules that START at halo pixel g, so the A operand of output
//     pixel m for taps dx = 5k .. 5k+4 is simply position m + 5k: 2 K-steps per dy, 18 MMAs per 128-pixel M-tile (the
//     pixel-pair granules of MODE_PAIR4 needed 27, the C8 input plane 45).  An N = 64 SS-mode MMA is bound by its operand
//     reads from shared memory (6 KB = 48 clk for 32 clk of math), so MMAs saved are time saved.
//   * the uint8 -> normalised 16-bit conversion is a 3 x 256-entry table in shared memory built in the prologue with the
//     reference's own three rounded fp32 steps (x / 255 - mean) / std — the operand bits equal those of the former
//     u8nchw_to_pair4 pass for every input value; pixels outside the image look up a table of zeros (conv0's zero padding
//     applies to the NORMALISED image).
// Tiling as conv_tc_kernel's column super-tiles: 8 px wide x 64 rows = 4 M-tiles stacked in y sharing one 72-row halo;
// warps: 0, 2, 3 producers (one halo row per thread: 12 32-bit loads, 48 table look-ups, 26 16-byte stores), 1 MMA issuer,
// 4..11 two epilogue sets (TMEM -> + bias (constant bank) -> LeakyReLU -> 16-bit C8 planes).
#pragma once
#include "srb_conv.cuh"

namespace srb {

enum : int { HEAD_NCHW_RGB = 0, HEAD_NHWC_RGB = 1, HEAD_NHWC_BGR = 2 };

struct HeadParams {
    int N, H, W;
    int tiles_x, tiles_y, num_tiles;
    const uint8_t* in;        // uint8 frames, `layout`
    int layout;
    int fast;                 // W % 4 == 0 and a 4-byte aligned base: 32-bit loads (a 4-pixel word is inside or outside the image)
    const uint8_t* wpack;     // [18 K-steps][2 K-halves][64 outputs][8 k] 16-bit
    float bias[64];           // constant bank
    float slope;
    float mean[3], stdv[3], maxv;
    void* y;                  // C8 planes [N][8][H][W][8]
};

#ifndef HD_STAGES
#define HD_STAGES 4
#endif
struct HDGeom {
    static constexpr int TW = 8, TH = 16, NT = 4, SUPER_H = TH * NT;       // 8 x 64 output pixels per pipeline step
    static constexpr int HALO_H = SUPER_H + 8;                             // 72 rows
    static constexpr int NPOS = 13;                                        // start positions m + 5k, m < 8, k < 2
    static constexpr int ROWP = NPOS * 16;                                 // 208 B between halo rows of one plane
    static constexpr int PLANE = HALO_H * ROWP;                            // 14 976 B between plane A and plane B
    static constexpr int STAGE_BYTES = 2 * PLANE;
    static constexpr int STAGE_STRIDE = (STAGE_BYTES + 1023) / 1024 * 1024;
    static constexpr int STAGES = HD_STAGES;
    static constexpr int KSTEPS = 18;                                      // 9 dy x 2
    static constexpr int WSTEP = 2 * 64 * 16;                              // 2 KB per K-step
    static constexpr int W_BYTES = KSTEPS * WSTEP;                         // 36 864
    static constexpr int LUT_BYTES = 2 * 3 * 256 * 2;                      // values + zeros
    static constexpr int OFF_W = 0, OFF_A = W_BYTES, OFF_LUT = OFF_A + STAGES * STAGE_STRIDE, OFF_BAR = OFF_LUT + LUT_BYTES;
    static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 256;
    static constexpr int THREADS = 384;
    static constexpr int PRODUCERS = 3;                                    // warps 0, 2, 3
    static_assert(SMEM_BYTES <= 227 * 1024 && STAGES <= 8, "shared memory budget / barrier layout");
    static_assert(HALO_H <= 32 * PRODUCERS, "one halo row per producer thread");
};

__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
    uint32_t r;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"(a));
    return r;
}

template <bool BF16>
__global__ void __launch_bounds__(HDGeom::THREADS, 1) head_ingest_kernel(const HeadParams p) {
    using G = HDGeom;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* s_w = smem + G::OFF_W;
    uint8_t* s_a = smem + G::OFF_A;
    uint16_t* s_lut = reinterpret_cast<uint16_t*>(smem + G::OFF_LUT);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
    uint64_t* full = bars;              // [STAGES] producers -> MMA
    uint64_t* empty = bars + 8;         // [STAGES] MMA -> producers
    uint64_t* tfull = bars + 16;        // [2] accumulator set complete      MMA -> epilogue
    uint64_t* tempty = bars + 18;       // [2] drained                       epilogue -> MMA
    uint64_t* wbar = bars + 20;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 22);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // look-up tables: [c][v] = 16-bit operand of ((v / maxv) - mean[c]) / std[c] (three rounded fp32 steps like the reference's
    // Normalize), followed by a table of zeros of the same shape for pixels outside the image
    for (int i = threadIdx.x; i < 3 * 256; i += G::THREADS) {
        const int c = i >> 8, v = i & 255;
        const float f = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, p.maxv), p.mean[c]), p.stdv[c]);
        s_lut[i] = (uint16_t)(pack2<BF16>(f, 0.f) & 0xffffu);
        s_lut[3 * 256 + i] = 0;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < G::STAGES; ++s) { mbar_init(&full[s], G::PRODUCERS); mbar_init(&empty[s], 1); }
        mbar_init(&tfull[0], 1); mbar_init(&tfull[1], 1);
        mbar_init(&tempty[0], 4 * EPI_SETS); mbar_init(&tempty[1], 4 * EPI_SETS);
        mbar_init(wbar, 1);
        fence_mbar_init();
    }
    if (warp == 0) { __syncwarp(); tmem_alloc(s_tmem, 512); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    pdl_launch_dependents();

    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const size_t HW = (size_t)p.H * p.W;

    if (warp == 0 || warp == 2 || warp == 3) {
        // ===================== producers: one halo row of the super-tile per thread =====================
        if (warp == 0 && elect_one()) {
            mbar_arrive_expect_tx(wbar, G::W_BYTES);
            for (int off = 0; off < G::W_BYTES; off += 12288) bulk_load_1d(s_w + off, p.wpack + off, 12288, wbar);
        }
        __syncwarp();
        pdl_wait();
        const int hr = (warp == 0 ? 0 : warp - 1) * 32 + lane;          // halo row 0 .. 71 (the rest idle)
        const bool active = hr < G::HALO_H;
        const uint32_t lut_u32 = smem_u32(s_lut), zlut_u32 = lut_u32 + 3 * 256 * 2;
        const bool nhwc = p.layout != HEAD_NCHW_RGB, bgr = p.layout == HEAD_NHWC_BGR;

        // q[c][k]: the bytes of channel c for halo pixels 4k .. 4k+3 (pixel h <-> image x = tx * 8 - 4 + h); vm: pixels inside the image
        auto fetch = [&](int tile, uint32_t (&q)[3][4], uint32_t& vm) {
            vm = 0;
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int k = 0; k < 4; ++k) q[c][k] = 0;
            if (tile >= p.num_tiles || !active) return;
            const int n = tile / tiles_per_img;
            const int r = tile - n * tiles_per_img;
            const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
            const int y = ty * G::SUPER_H - 4 + hr;
            if (y < 0 || y >= p.H) return;
            const int xb = tx * G::TW - 4;
            if (p.fast) {
                if (!nhwc) {
                    const uint8_t* rowp = p.in + ((size_t)n * 3 * p.H + y) * p.W;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int x = xb + 4 * k;
                        if (x >= 0 && x < p.W) {
                            vm |= 0xFu << (4 * k);
#pragma unroll
                            for (int c = 0; c < 3; ++c) q[c][k] = __ldg(reinterpret_cast<const uint32_t*>(rowp + (size_t)c * HW + x));
                        }
                    }
                } else {
                    const uint8_t* rowp = p.in + ((size_t)n * p.H + y) * p.W * 3;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int x = xb + 4 * k;
                        if (x >= 0 && x < p.W) {
                            vm |= 0xFu << (4 * k);
                            const uint32_t* wp = reinterpret_cast<const uint32_t*>(rowp + (size_t)x * 3);
                            const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2);   // 4 pixels: c0c1c2 c0c1c2 c0c1c2 c0c1c2
                            const uint32_t a0 = __byte_perm(__byte_perm(w0, w1, 0x0630), w2, 0x5210);   // bytes 0, 3, 6, 9
                            const uint32_t a1 = __byte_perm(__byte_perm(w0, w1, 0x0741), w2, 0x6210);   // bytes 1, 4, 7, 10
                            // bytes 2, 5, 8, 11: w0.b2, w1.b1, w2.b0, w2.b3
                            const uint32_t a2 = __byte_perm(__byte_perm(w0, w1, 0x0052), w2, 0x7410);
                            q[0][k] = bgr ? a2 : a0; q[1][k] = a1; q[2][k] = bgr ? a0 : a2;
                        }
                    }
                }
            } else {
                // any width / alignment: byte loads with a bounds test per pixel
#pragma unroll
                for (int k = 0; k < 4; ++k)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int x = xb + 4 * k + j;
                        if (x >= 0 && x < p.W) {
                            vm |= 1u << (4 * k + j);
#pragma unroll
                            for (int c = 0; c < 3; ++c) {
                                const int mc = bgr ? 2 - c : c;
                                const uint8_t v = nhwc ? __ldg(p.in + (((size_t)n * p.H + y) * p.W + x) * 3 + mc)
                                                       : __ldg(p.in + ((size_t)n * 3 * p.H + y) * p.W + (size_t)c * HW + x);
                                q[c][k] |= (uint32_t)v << (8 * j);
                            }
                        }
                    }
            }
        };

        uint32_t q[3][4], vm;
        int tile = blockIdx.x;
        fetch(tile, q, vm);
        int s = 0; uint32_t ph = 0;
#pragma unroll 1
        for (; tile < p.num_tiles; tile += gridDim.x) {
            // the next super-tile's bytes are requested before this one is converted: their latency hides behind the table look-ups
            uint32_t qn[3][4], vmn;
            fetch(tile + gridDim.x, qn, vmn);
            mbar_wait(&empty[s], ph ^ 1, 71);
            if (active) {
                // 16-bit operands of the 16 halo pixels: h01 = (c0, c1), hx = (c2, next pixel's c0), h12 = (c1, c2), hz = (c2, 0)
                uint32_t c0h[17], h01[17], hx[17];
#pragma unroll
                for (int h = 0; h < 16; ++h) {
                    const uint32_t tb = (vm >> h) & 1u ? lut_u32 : zlut_u32;
                    const uint32_t b0 = (q[0][h >> 2] >> (8 * (h & 3))) & 255u, b1 = (q[1][h >> 2] >> (8 * (h & 3))) & 255u,
                                   b2 = (q[2][h >> 2] >> (8 * (h & 3))) & 255u;
                    c0h[h] = lds_u16(tb + b0 * 2);
                    const uint32_t c1 = lds_u16(tb + 512 + b1 * 2), c2 = lds_u16(tb + 1024 + b2 * 2);
                    h01[h] = c0h[h] | (c1 << 16);
                    hx[h] = c2;                                            // (the neighbour's c0 is added below)
                }
                c0h[16] = 0; h01[16] = 0; hx[16] = 0;
                uint32_t h12[17], hz[17];
#pragma unroll
                for (int h = 0; h < 17; ++h) {
                    hz[h] = hx[h];
                    h12[h] = (h01[h] >> 16) | (hx[h] << 16);
                }
#pragma unroll
                for (int h = 0; h < 16; ++h) hx[h] |= c0h[h + 1] << 16;
                const uint32_t rowa = smem_u32(s_a + s * G::STAGE_STRIDE) + (uint32_t)hr * G::ROWP;
#pragma unroll
                for (int g = 0; g < G::NPOS; ++g) {
                    sts_u4x(rowa + g * 16, make_uint4(h01[g], hx[g], h12[g + 1], h01[g + 2]));
                    sts_u4x(rowa + G::PLANE + g * 16, make_uint4(hx[g + 2], h12[g + 3], h01[g + 4], hz[g + 4]));
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[s]);
#pragma unroll
            for (int c = 0; c < 3; ++c)
#pragma unroll
                for (int k = 0; k < 4; ++k) q[c][k] = qn[c][k];
            vm = vmn;
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: 4 M-tiles x 18 K-steps per super-tile =====================
        constexpr uint32_t idesc = umma_idesc_f16(128, 64, BF16 ? 1 : 0);
        const uint64_t a_desc0 = umma_desc_nosw(0, G::PLANE, G::ROWP);
        const uint64_t b_desc0 = umma_desc_nosw(smem_u32(s_w), 64 * 16, 128);
        mbar_wait(wbar, 0, 72);
        int s = 0; uint32_t ph = 0; int it = 0;
#pragma unroll 1
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int b = it & 1;
            mbar_wait(&tempty[b], (uint32_t)(((it >> 1) & 1) ^ 1), 73);
            mbar_wait(&full[s], ph, 74);
            tc_fence_after();
            if (elect_one()) {
                const uint64_t a_desc = a_desc0 + (uint64_t)(smem_u32(s_a + s * G::STAGE_STRIDE) >> 4);
                const uint32_t d_tmem = tmem_base + b * (G::NT * 64);
#pragma unroll 1
                for (int dy = 0; dy < 9; ++dy)
#pragma unroll
                    for (int k = 0; k < 2; ++k) {
                        const uint32_t ao = (uint32_t)(dy * G::ROWP + 5 * k * 16) >> 4;
                        const uint64_t bd = b_desc0 + (uint64_t)(((dy * 2 + k) * G::WSTEP) >> 4);
#pragma unroll
                        for (int t = 0; t < G::NT; ++t)
                            umma_f16_ss(d_tmem + t * 64, a_desc + ao + t * ((G::TH * G::ROWP) >> 4), bd, idesc, (dy | k) != 0);
                    }
                umma_commit(&empty[s]);
                umma_commit(&tfull[b]);
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
        }
    } else if (warp >= 4) {
        // ===================== epilogue: two sets of 4 warps; set e drains M-tiles e, e + 2 =====================
        const int lg = warp & 3;
        const int eset = (warp - 4) >> 2;
        const int L = lg * 32 + lane;
        pdl_wait();                     // x0 may still be read (long skip) by the previous forward's conv1
        int it = 0;
#pragma unroll 1
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int b = it & 1;
            const int n = tile / tiles_per_img;
            const int r = tile - n * tiles_per_img;
            const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
            const uint32_t t_addr0 = tmem_base + b * (G::NT * 64) + ((uint32_t)(lg * 32) << 16);
            mbar_wait<20>(&tfull[b], (uint32_t)((it >> 1) & 1), 75);
            tc_fence_after();
#pragma unroll 1
            for (int t = eset; t < G::NT; t += EPI_SETS) {
                const int y = ty * G::SUPER_H + t * G::TH + (L >> 3), x = tx * G::TW + (L & 7);
                const bool valid = y < p.H && x < p.W;
                uint4* yp = reinterpret_cast<uint4*>(p.y) + ((size_t)n * 8 * HW + (size_t)y * p.W + x);
                auto half = [&](auto off) {
                    constexpr int OFF = decltype(off)::value;
                    uint32_t rr[32];
                    tmem_ld32(t_addr0 + t * 64 + OFF, rr);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint32_t o[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            float a = __fadd_rn(__uint_as_float(rr[8 * j + 2 * k]), p.bias[OFF + 8 * j + 2 * k]);
                            float c = __fadd_rn(__uint_as_float(rr[8 * j + 2 * k + 1]), p.bias[OFF + 8 * j + 2 * k + 1]);
                            a = fmaxf(a, __fmul_rn(a, p.slope)); c = fmaxf(c, __fmul_rn(c, p.slope));
                            o[k] = pack2<BF16>(a, c);
                        }
                        if (valid) yp[(size_t)(OFF / 8 + j) * HW] = make_uint4(o[0], o[1], o[2], o[3]);
                    }
                };
                half(std::integral_constant<int, 0>{});
                half(std::integral_constant<int, 32>{});
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[b]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace srb
