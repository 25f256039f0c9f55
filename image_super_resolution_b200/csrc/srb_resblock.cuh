// Fused residual block of the model.pt trunk:  y = x + convB(LeakyReLU(convA(x)))   (ResidualBlock1.forward,
// utils/models.py:208-209; 16 of the graph's 19 convs and 81 % of its FLOPs — SURVEY §8(a) a3).
//
// Why a new kernel (measured): an SS-mode tcgen05.mma re-reads its 128-row A tile from shared memory at
// 128 B/clk, so N = 32 / 64 MMAs run at 40 % / 67 % of the tensor pipe, and an unfused layer chain moves 512 B/px of
// HBM traffic per block.  This kernel removes both limits:
//   * ROW FOLD.  An M-tile is ONE image row of 128 pixels (16 groups of 8 px; a group is any 8-pixel column strip of
//     any image, so two 64-wide images fill a tile).  The three dy taps are folded into N: one MMA on input row r with
//     B = [W(dy=2) | W(dy=1) | W(dy=0)] (N = 3*Cout) accumulates into the THREE TMEM accumulators of output rows
//     r-1, r, r+1, which are adjacent column ranges — the sum over dy happens inside TMEM, for free.  N = 96 / 192
//     cuts the A re-reads 3x (86 % / 100 % of the pipe by the measured cost model).  dx taps stay address shifts.
//   * MARCHING.  A CTA walks down the rows of its strips.  x row r is loaded once (TMA, one 10-pixel box per group:
//     private 1-px halo, zero fill at image borders), feeds convA; the finished hidden row is written by an epilogue
//     warp-group into shared memory (never to HBM) as convB's A operand; convB's finished output row gets bias +
//     residual and is stored.  HBM traffic per block: read x once, write y once (256 B/px).
//   * Accumulator rings in TMEM (hidden: RH x 32 columns, output: RY x 64 columns = 512), drained and re-zeroed by the
//     epilogues, so every MMA accumulates; an MMA whose three rows wrap around the ring is split in two.
// Warp roles (512 threads): 0 TMA producer, 1 MMA issuer, 4-7 hidden epilogue, 8-15 output epilogue (two sets that
// alternate rows).  convA and convB MMAs are interleaved (A A B) so consecutive MMAs hit different accumulators.
#pragma once
#include <cuda.h>
#include "srb_ptx.cuh"
#include "srb_conv.cuh"

namespace srb {

struct ResblockParams {
    int N, H, W, GX;          // GX = groups (8 px) per image row; 16 % GX == 0
    int total_rows;           // strip-sets x H: the flattened (strip-set, output row) space split evenly over CTAs
    const uint8_t* wA;        // fold-packed 64 -> 32:  12 K-steps x [2][96][8]
    const uint8_t* wB;        // fold-packed 32 -> 64:   6 K-steps x [2][192][8]
    const float* biasA;       // 32
    const float* biasB;       // 64
    const void* x;            // [N][8][H][W][8]
    void* y;                  // [N][8][H][W][8]
    float slope;
};

struct RBGeom {
    static constexpr int CA = 64, CH = 32, CB = 64;            // channels: in, hidden, out
    static constexpr int XBOX = 8 * 10 * 16;                    // one group box of x: [8 planes][10 px][16 B]
    static constexpr int XROW = 16 * XBOX;                      // 20480 B per x row
    static constexpr int XSTAGES = 4;
    static constexpr int HBOX = 4 * 10 * 16;                    // hidden group box: [4 planes][10 px][16 B]
    static constexpr int HROW = 16 * HBOX;                      // 10240 B
    static constexpr int HSLOTS = 3;
    static constexpr int KA = 3 * (CA / 16), KB = 3 * (CH / 16);           // 12, 6 K-steps per row
    static constexpr int WA_STEP = 2 * 3 * CH * 16, WB_STEP = 2 * 3 * CB * 16;   // 3072, 6144
    static constexpr int WA_BYTES = KA * WA_STEP, WB_BYTES = KB * WB_STEP;       // 36864 each
    static constexpr int RH = 6, RY = 5;                        // TMEM accumulator rings
    static constexpr int TM_H = 0, TM_Y = RH * CH;              // column bases: 0, 192 (+ 5*64 = 512)
    static constexpr int OFF_WA = 0, OFF_WB = WA_BYTES, OFF_X = OFF_WB + WB_BYTES, OFF_H = OFF_X + XSTAGES * XROW;
    static constexpr int OFF_BIAS = OFF_H + HSLOTS * HROW, OFF_BAR = OFF_BIAS + (CH + CB) * 4;
    static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 512;
    static constexpr int THREADS = 512;
    static_assert(RH * CH + RY * CB == 512, "the two accumulator rings fill TMEM");
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
        "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]),
          "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]),
          "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_zero32(uint32_t taddr) {
    uint32_t z[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) z[i] = 0u;
    tmem_st32(taddr, z);
}
__device__ __forceinline__ void sts_u4(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// a segment: output rows [y0, y1) of strip-set ss, with the hidden / input row ranges it needs (clipped to the image)
struct RBSeg { int ss, y0, y1, hlo, hhi, xlo, xhi; };
__device__ __forceinline__ bool rb_next_seg(const ResblockParams& p, int& g, int g1, RBSeg& s) {
    if (g >= g1) return false;
    s.ss = g / p.H;
    s.y0 = g - s.ss * p.H;
    int n = g1 - g;
    s.y1 = s.y0 + n < p.H ? s.y0 + n : p.H;
    s.hlo = s.y0 > 0 ? s.y0 - 1 : 0;
    s.hhi = s.y1 < p.H ? s.y1 : p.H - 1;
    s.xlo = s.hlo > 0 ? s.hlo - 1 : 0;
    s.xhi = s.hhi + 1 < p.H ? s.hhi + 1 : p.H - 1;
    g += s.y1 - s.y0;
    return true;
}

// Issue one folded K-step: D slots of rows [c_lo, c_hi] (ring counters, consecutive) += A * B[slice s0 ..], splitting
// at the ring's wrap-around.  CO = output channels per row slot.
template <int CO, int R, bool BF16>
__device__ __forceinline__ void rb_mma_fold(uint32_t tm_ring, uint64_t a_desc, uint64_t b_desc, int c_lo, int nsl, int s0) {
    const int slot = c_lo % R;
    int n1 = R - slot < nsl ? R - slot : nsl;
    uint32_t idesc = umma_idesc_f16(128, n1 * CO, BF16 ? 1 : 0);
    umma_f16_ss(tm_ring + slot * CO, a_desc, b_desc + (uint64_t)((s0 * CO * 16) >> 4), idesc, 1u);
    if (n1 < nsl) {
        idesc = umma_idesc_f16(128, (nsl - n1) * CO, BF16 ? 1 : 0);
        umma_f16_ss(tm_ring, a_desc, b_desc + (uint64_t)(((s0 + n1) * CO * 16) >> 4), idesc, 1u);
    }
}

template <bool BF16>
__global__ void __launch_bounds__(RBGeom::THREADS, 1) resblock_kernel(const __grid_constant__ CUtensorMap tmap_x,
                                                                      const ResblockParams p) {
    using G = RBGeom;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* s_wA = smem + G::OFF_WA;
    uint8_t* s_wB = smem + G::OFF_WB;
    uint8_t* s_x = smem + G::OFF_X;
    uint8_t* s_h = smem + G::OFF_H;
    float* s_bias = reinterpret_cast<float*>(smem + G::OFF_BIAS);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
    uint64_t* xfull = bars;            // [4]  TMA -> MMA
    uint64_t* xempty = bars + 4;       // [4]  MMA -> TMA
    uint64_t* hafull = bars + 8;       // [RH] hidden accumulator row complete        MMA -> hidden epilogue
    uint64_t* haempty = bars + 14;     // [RH] drained and re-zeroed                  hidden epilogue -> MMA
    uint64_t* hsfull = bars + 20;      // [3]  hidden row written to shared memory    hidden epilogue -> MMA
    uint64_t* hsempty = bars + 23;     // [3]  consumed by convB                      MMA -> hidden epilogue
    uint64_t* yafull = bars + 26;      // [RY] output accumulator row complete        MMA -> output epilogue
    uint64_t* yaempty = bars + 31;     // [RY] drained and re-zeroed                  output epilogue -> MMA
    uint64_t* wbar = bars + 36;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 40);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // zero the hidden ring once: halo slots at image borders are never written and must read as zero padding
    for (int i = threadIdx.x; i < G::HSLOTS * G::HROW / 16; i += G::THREADS)
        reinterpret_cast<uint4*>(s_h)[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < G::CH + G::CB; i += G::THREADS) s_bias[i] = i < G::CH ? p.biasA[i] : p.biasB[i - G::CH];
    if (threadIdx.x == 0) {
        for (int i = 0; i < G::XSTAGES; ++i) { mbar_init(&xfull[i], 1); mbar_init(&xempty[i], 1); }
        for (int i = 0; i < G::RH; ++i) { mbar_init(&hafull[i], 1); mbar_init(&haempty[i], 4); }
        for (int i = 0; i < G::HSLOTS; ++i) { mbar_init(&hsfull[i], 4); mbar_init(&hsempty[i], 1); }
        for (int i = 0; i < G::RY; ++i) { mbar_init(&yafull[i], 1); mbar_init(&yaempty[i], 4); }
        mbar_init(wbar, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tmap_x);
    }
    fence_proxy_async_smem();
    if (warp == 0) { __syncwarp(); tmem_alloc(s_tmem, 512); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    if (warp >= 4 && warp < 8) {       // all accumulators start at zero: every MMA of this kernel accumulates
        const uint32_t lg = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll 1
        for (int c = 0; c < 512; c += 32) tmem_zero32(tmem_base + lg + c);
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();

    const int g_begin = (int)((long long)blockIdx.x * p.total_rows / gridDim.x);
    const int g_end = (int)((long long)(blockIdx.x + 1) * p.total_rows / gridDim.x);
    const size_t HW = (size_t)p.H * p.W;

    if (warp == 0) {
        // ===================== TMA producer: filters once, then one x row (16 group boxes) per stage =====================
        if (elect_one()) {
            mbar_arrive_expect_tx(wbar, G::WA_BYTES + G::WB_BYTES);
            for (int off = 0; off < G::WA_BYTES; off += 12288) bulk_load_1d(s_wA + off, p.wA + off, 12288, wbar);
            for (int off = 0; off < G::WB_BYTES; off += 12288) bulk_load_1d(s_wB + off, p.wB + off, 12288, wbar);
        }
        __syncwarp();
        int xc = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg)) {
            const int q = sg.ss * 16 + (lane & 15);
            const int n = q / p.GX, gx = q - n * p.GX;
#pragma unroll 1
            for (int r = sg.xlo; r <= sg.xhi; ++r, ++xc) {
                const int st = xc % G::XSTAGES;
                mbar_wait(&xempty[st], (uint32_t)(((xc / G::XSTAGES) & 1) ^ 1), 31);
                if (lane == 0) mbar_arrive_expect_tx(&xfull[st], G::XROW);
                __syncwarp();
                if (lane < 16)
                    tma_load_4d(s_x + st * G::XROW + lane * G::XBOX, &tmap_x, &xfull[st], (gx * 8 - 1) * 4, r, 0, n);
                __syncwarp();
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        const uint64_t adescA0 = umma_desc_nosw(0, 160, G::XBOX);           // x row: planes 160 B apart, groups XBOX apart
        const uint64_t adescB0 = umma_desc_nosw(0, 160, G::HBOX);           // hidden row
        const uint64_t bdescA0 = umma_desc_nosw(smem_u32(s_wA), 3 * G::CH * 16, 128);
        const uint64_t bdescB0 = umma_desc_nosw(smem_u32(s_wB), 3 * G::CB * 16, 128);
        mbar_wait(wbar, 0, 32);
        int xc = 0, hbase = 0, ybase = 0, h_acq = 0, y_acq = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg)) {
            const int nx = sg.xhi - sg.xlo + 1, nh = sg.hhi - sg.hlo + 1, ny = sg.y1 - sg.y0;
            const int lag = sg.hlo - sg.xlo + 3;
#pragma unroll 1
            for (int t = 0; t < nx || t - lag < nh; ++t) {
                const bool doA = t < nx;
                const bool doB = t - lag >= 0 && t - lag < nh;
                const int r = sg.xlo + t, h = sg.hlo + t - lag;
                int a_lo = 0, a_n = 0, a_s0 = 0, b_lo = 0, b_n = 0, b_s0 = 0, st = 0, hs = 0;
                if (doA) {
                    st = xc % G::XSTAGES;
                    mbar_wait(&xfull[st], (uint32_t)((xc / G::XSTAGES) & 1), 33);
                    const int lo = r - 1 > sg.hlo ? r - 1 : sg.hlo, hi = r + 1 < sg.hhi ? r + 1 : sg.hhi;
                    a_lo = hbase + lo - sg.hlo; a_n = hi - lo + 1; a_s0 = lo - (r - 1);
                    for (; h_acq < a_lo + a_n; ++h_acq)
                        mbar_wait(&haempty[h_acq % G::RH], (uint32_t)(((h_acq / G::RH) & 1) ^ 1), 34);
                }
                if (doB) {
                    const int hc = hbase + h - sg.hlo;
                    hs = hc % G::HSLOTS;
                    mbar_wait(&hsfull[hs], (uint32_t)((hc / G::HSLOTS) & 1), 35);
                    const int lo = h - 1 > sg.y0 ? h - 1 : sg.y0, hi = h + 1 < sg.y1 - 1 ? h + 1 : sg.y1 - 1;
                    b_lo = ybase + lo - sg.y0; b_n = hi - lo + 1; b_s0 = lo - (h - 1);
                    for (; y_acq < b_lo + b_n; ++y_acq)
                        mbar_wait(&yaempty[y_acq % G::RY], (uint32_t)(((y_acq / G::RY) & 1) ^ 1), 36);
                }
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t adA = adescA0 + (uint64_t)(smem_u32(s_x + st * G::XROW) >> 4);
                    const uint64_t adB = adescB0 + (uint64_t)(smem_u32(s_h + hs * G::HROW) >> 4);
#pragma unroll
                    for (int k = 0; k < G::KB; ++k) {
#pragma unroll
                        for (int kk = 2 * k; kk < 2 * k + 2; ++kk)
                            if (doA) {
                                const int dx = kk >> 2, c2 = kk & 3;
                                rb_mma_fold<G::CH, G::RH, BF16>(tmem_base + G::TM_H, adA + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4),
                                                                bdescA0 + (uint64_t)((kk * G::WA_STEP) >> 4), a_lo, a_n, a_s0);
                            }
                        if (doB) {
                            const int dx = k >> 1, c2 = k & 1;
                            rb_mma_fold<G::CB, G::RY, BF16>(tmem_base + G::TM_Y, adB + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4),
                                                            bdescB0 + (uint64_t)((k * G::WB_STEP) >> 4), b_lo, b_n, b_s0);
                        }
                    }
                    if (doA) {
                        umma_commit(&xempty[st]);
                        if (r - 1 >= sg.hlo) umma_commit(&hafull[(hbase + r - 1 - sg.hlo) % G::RH]);
                        if (r == sg.xhi && r <= sg.hhi) umma_commit(&hafull[(hbase + r - sg.hlo) % G::RH]);
                    }
                    if (doB) {
                        umma_commit(&hsempty[hs]);
                        if (h - 1 >= sg.y0) umma_commit(&yafull[(ybase + h - 1 - sg.y0) % G::RY]);
                        if (h == sg.hhi && h <= sg.y1 - 1) umma_commit(&yafull[(ybase + h - sg.y0) % G::RY]);
                    }
                }
                __syncwarp();
                if (doA) ++xc;
            }
            hbase += nh; ybase += ny;
        }
    }
    else if (warp >= 4 && warp < 8) {
        // ===================== hidden epilogue: TMEM -> bias, LeakyReLU -> 16-bit -> shared memory (convB's A operand) =====================
        const int lg = warp & 3;
        const int L = lg * 32 + lane;              // accumulator lane = pixel of the row
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + G::TM_H + ((uint32_t)(lg * 32) << 16);
        const uint32_t s_h_u32 = smem_u32(s_h);
        const uint32_t s_bias_u32 = smem_u32(s_bias);
        int hbase = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg)) {
            const int q = sg.ss * 16 + grp;
            const int n = q / p.GX, gx = q - n * p.GX;
            const bool valid = n < p.N && gx * 8 + i < p.W;
            const bool left = i == 0 && gx > 0, right = i == 7 && gx < p.GX - 1;
#pragma unroll 1
            for (int h = sg.hlo; h <= sg.hhi; ++h) {
                const int c = hbase + h - sg.hlo;
                const int slot = c % G::RH, hs = c % G::HSLOTS;
                mbar_wait(&hafull[slot], (uint32_t)((c / G::RH) & 1), 41);
                tc_fence_after();
                uint32_t rr[32];
                tmem_ld32(t_lane + slot * G::CH, rr);
                tmem_ld_wait();
                tmem_zero32(t_lane + slot * G::CH);
                mbar_wait(&hsempty[hs], (uint32_t)(((c / G::HSLOTS) & 1) ^ 1), 42);
                const uint32_t box = s_h_u32 + hs * G::HROW + grp * G::HBOX;
#pragma unroll
                for (int pl = 0; pl < 4; ++pl) {
                    const float4 b0 = lds_f4(s_bias_u32 + (8 * pl) * 4), b1 = lds_f4(s_bias_u32 + (8 * pl + 4) * 4);
                    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                    float v[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        float tt = __uint_as_float(rr[8 * pl + k]) + bb[k];
                        tt = fmaxf(tt, tt * p.slope);
                        v[k] = valid ? tt : 0.f;     // outside the image the hidden activation is convB's zero padding
                    }
                    uint4 o;
                    o.x = pack2<BF16>(v[0], v[1]); o.y = pack2<BF16>(v[2], v[3]);
                    o.z = pack2<BF16>(v[4], v[5]); o.w = pack2<BF16>(v[6], v[7]);
                    sts_u4(box + pl * 160 + (i + 1) * 16, o);
                    if (left) sts_u4(box - G::HBOX + pl * 160 + 9 * 16, o);      // right halo slot of the group to the left
                    if (right) sts_u4(box + G::HBOX + pl * 160, o);              // left halo slot of the group to the right
                }
                tmem_st_wait();
                tc_fence_before();
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) { mbar_arrive(&haempty[slot]); mbar_arrive(&hsfull[hs]); }
            }
            hbase += sg.hhi - sg.hlo + 1;
        }
    } else if (warp >= 8) {
        // ===================== output epilogue: two sets of 4 warps alternate rows; + bias + x (residual) -> y =====================
        const int lg = warp & 3;
        const int set = (warp - 8) >> 2;
        const int L = lg * 32 + lane;
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + G::TM_Y + ((uint32_t)(lg * 32) << 16);
        const uint32_t s_bias_u32 = smem_u32(s_bias) + G::CH * 4;
        int ybase = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg)) {
            const int q = sg.ss * 16 + grp;
            const int n = q / p.GX, gx = q - n * p.GX;
            const int xx = gx * 8 + i;
            const bool valid = n < p.N && xx < p.W;
#pragma unroll 1
            for (int y = sg.y0; y < sg.y1; ++y) {
                const int c = ybase + y - sg.y0;
                if ((c & 1) != set) continue;
                const int slot = c % G::RY;
                const size_t pix = valid ? ((size_t)n * 8 * HW + (size_t)y * p.W + xx) : 0;
                uint4 res[8];
#pragma unroll
                for (int pl = 0; pl < 8; ++pl)
                    res[pl] = valid ? ldg_nc16(reinterpret_cast<const uint8_t*>(p.x) + (pix + pl * HW) * 16) : make_uint4(0, 0, 0, 0);
                mbar_wait(&yafull[slot], (uint32_t)((c / G::RY) & 1), 43);
                tc_fence_after();
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t rr[32];
                    tmem_ld32(t_lane + slot * G::CB + half * 32, rr);
                    tmem_ld_wait();
                    tmem_zero32(t_lane + slot * G::CB + half * 32);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int pl = half * 4 + j;
                        const float4 b0 = lds_f4(s_bias_u32 + (8 * pl) * 4), b1 = lds_f4(s_bias_u32 + (8 * pl + 4) * 4);
                        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                        const uint32_t qq[4] = {res[pl].x, res[pl].y, res[pl].z, res[pl].w};
                        float v[8];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpack2<BF16>(qq[k]);
                            v[2 * k] = __uint_as_float(rr[8 * j + 2 * k]) + bb[2 * k] + f.x;
                            v[2 * k + 1] = __uint_as_float(rr[8 * j + 2 * k + 1]) + bb[2 * k + 1] + f.y;
                        }
                        if (valid) {
                            uint4 o;
                            o.x = pack2<BF16>(v[0], v[1]); o.y = pack2<BF16>(v[2], v[3]);
                            o.z = pack2<BF16>(v[4], v[5]); o.w = pack2<BF16>(v[6], v[7]);
                            *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.y) + (pix + pl * HW) * 16) = o;
                        }
                    }
                }
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&yaempty[slot]);
            }
            ybase += sg.y1 - sg.y0;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace srb
