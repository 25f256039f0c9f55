// Row-fold 3x3 conv 64 -> 64 (+ bias, optional LeakyReLU, optional residual) on tcgen05 — the GEMM mapping of the
// fused residual block (srb_resblock.cuh) for a single conv: conv1 of the model.pt graph (SURVEY §8 a4,
// utils/models.py:690 + the long skip at :704).
//
// M-tile = one row of 128 pixels (16 groups x 8 px, each its own 10-px TMA box), the three dy taps folded into
// N = 3 x 64 = 192 with the MMA's D operand spanning three adjacent row accumulators in TMEM (input row r adds into
// output rows r-1, r, r+1), K-steps = 3 dx x 4 channel pairs.  Against the tiled kernel (36 N=64 MMAs per 128 px, each
// re-reading its 4 KB A tile: 216 KB of operand reads) this is 12 N=192 MMAs, 120 KB — compute-bound instead of
// shared-memory-bound.  A CTA marches down a contiguous run of rows; accumulator rows live in a ring of 8 x 64 TMEM
// columns and are re-zeroed by the epilogue that drains them (the bias is a constant-bank operand of the epilogue's adds).
#pragma once
#include "srb_resblock.cuh"

namespace srb {

struct RowconvParams {
    int N, H, W, GX;
    int TR, TS, OV;           // tile geometry (rb_slot); OV = 0: every group carries its own halo
    int total_rows;
    const uint8_t* w;         // packed [12 K-steps][2 granules][192 rows][8 k] 16-bit
    float bias[64];
    const void* res;          // C8 planes [N][8][H][W][8] added to the result, or nullptr
    void* y;
    float slope;              // LeakyReLU slope applied before the residual add when lrelu != 0
    int lrelu;
    int gm;                   // narrow rows in group-major slot order: one TMA box [10 px][8 planes][TR images] per group column
    int dbg_skip;
};
template <class Params>
__device__ __forceinline__ RBSlot rc_slot(const Params& p, int ss, int j) {
    if (!p.gm) return rb_slot(p, ss, j);
    RBSlot s;
    s.gx = j / p.TR;
    s.n = ss * p.TR + (j - s.gx * p.TR);
    s.in = s.gx < p.GX && s.n < p.N;
    s.out = s.in;
    if (!s.in) { s.n = p.N; s.gx = 0; }
    return s;
}

#ifndef RC_XSTAGES
#define RC_XSTAGES 4
#endif
struct RCGeom {
    static constexpr int C = 64;
    static constexpr int XBOX = 8 * 10 * 16, XROW = 16 * XBOX, XSTAGES = RC_XSTAGES;
    static constexpr int KS = 3 * (C / 16);                   // 12 K-steps per row
    static constexpr int W_STEP = 2 * 3 * C * 16;              // 6144
    static constexpr int W_BYTES = KS * W_STEP;                // 73728
    static constexpr int RY = 8;                               // accumulator ring: 8 x 64 = 512 TMEM columns
    static constexpr int OFF_W = 0, OFF_X = W_BYTES, OFF_BIAS = OFF_X + XSTAGES * XROW, OFF_BAR = OFF_BIAS + C * 4;
    static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 512;
    static_assert(XSTAGES <= 8, "barrier layout");
    static constexpr int THREADS = 640;                        // 4 control warps + two epilogue sets of 8 warps
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

template <bool BF16>
__global__ void __launch_bounds__(RCGeom::THREADS, 1) rowconv_kernel(const __grid_constant__ CUtensorMap tmap_x,
                                                                     const RowconvParams p) {
    using G = RCGeom;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);     // keeps the shared address space
    uint8_t* s_w = smem + G::OFF_W;
    uint8_t* s_x = smem + G::OFF_X;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
    uint64_t* xfull = bars;            // [XSTAGES] TMA -> MMA
    uint64_t* xempty = bars + 8;       // [XSTAGES] MMA -> TMA
    uint64_t* yafull = bars + 16;      // [RY] accumulator row complete      MMA -> epilogue
    uint64_t* yaempty = bars + 24;     // [RY] drained and re-armed          epilogue -> MMA
    uint64_t* wbar = bars + 32;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 33);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (p.gm && p.GX * p.TR < 16)      // slots no TMA box ever fills: their operand rows read as zeros
        for (int i = threadIdx.x; i < G::XSTAGES * G::XROW / 16; i += G::THREADS) reinterpret_cast<uint4*>(s_x)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) {
        for (int i = 0; i < G::XSTAGES; ++i) { mbar_init(&xfull[i], 1); mbar_init(&xempty[i], 1); }
        for (int i = 0; i < G::RY; ++i) { mbar_init(&yafull[i], 1); mbar_init(&yaempty[i], 8); }
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
    if (warp >= 4 && warp < 8) {       // every MMA accumulates: all rows start at zero (the epilogue adds the bias)
        const uint32_t lg = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll 1
        for (int c = 0; c < 512; c += 32) {
            tmem_zero32(tmem_base + lg + c);
        }
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    pdl_launch_dependents();

    const int g_begin = (int)((long long)blockIdx.x * p.total_rows / gridDim.x);
    const int g_end = (int)((long long)(blockIdx.x + 1) * p.total_rows / gridDim.x);
    const size_t HW = (size_t)p.H * p.W;

    if (warp == 0) {
        // ===================== TMA producer: filters once, then one x row (16 group boxes) per stage =====================
        if (elect_one()) {
            mbar_arrive_expect_tx(wbar, G::W_BYTES);
            for (int off = 0; off < G::W_BYTES; off += 12288) bulk_load_1d(s_w + off, p.w + off, 12288, wbar);
        }
        __syncwarp();
        pdl_wait();
        int xc = 0, g = g_begin;
        const int gm_box = p.TR * G::XBOX;
        const uint32_t xbytes = p.gm ? (uint32_t)(p.GX * gm_box) : (uint32_t)G::XROW;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg)) {
            const RBSlot sl = rc_slot(p, sg.ss, lane & 15);
            const int n = sl.n, gx = sl.gx;
            const int gm_n0 = sg.ss * p.TR < p.N ? sg.ss * p.TR : p.N;
#pragma unroll 1
            for (int r = sg.hlo; r <= sg.hhi; ++r, ++xc) {          // x rows y0-1 .. y1 (clipped): RBSeg's hidden-row range
                const int st = xc % G::XSTAGES;
                mbar_wait<RB_CTRL>(&xempty[st], (uint32_t)(((xc / G::XSTAGES) & 1) ^ 1), 51);
                if SRB_DBG(p.dbg_skip & 1) { if (lane == 0) mbar_arrive(&xfull[st]); __syncwarp(); continue; }
                if (lane == 0) mbar_arrive_expect_tx(&xfull[st], xbytes);
                __syncwarp();
                if (p.gm) {         // lane gx: the box of group column gx for the tile's TR images (images beyond N: zero-filled)
                    if (lane < p.GX)
                        tma_load_4d(s_x + st * G::XROW + lane * gm_box, &tmap_x, &xfull[st], (lane * 8 - 1) * 4, r, 0, gm_n0);
                } else if (lane < 16)
                    tma_load_4d(s_x + st * G::XROW + lane * G::XBOX, &tmap_x, &xfull[st], (gx * 8 - 1) * 4, r, 0, n);
                __syncwarp();
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: one x row per step, N = 3 x 64 into the accumulator ring =====================
        const uint64_t adesc0 = umma_desc_nosw(0, 160, G::XBOX);
        const uint64_t bdesc0 = umma_desc_nosw(smem_u32(s_w), 3 * G::C * 16, 128);
        mbar_wait(wbar, 0, 52);
        int xst = 0; uint32_t xph = 0;
        int yslot0 = 0;                                // ring slot of the segment's first output row
        int acq_slot = 0; uint32_t acq_ph = 1;
        int ybase = 0, y_acq = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg)) {
#pragma unroll 1
            for (int r = sg.hlo; r <= sg.hhi; ++r) {
                mbar_wait<RB_CTRL>(&xfull[xst], xph, 53);
                const int lo = r - 1 > sg.y0 ? r - 1 : sg.y0, hi = r + 1 < sg.y1 - 1 ? r + 1 : sg.y1 - 1;
                const int b_lo = ybase + lo - sg.y0, b_n = hi - lo + 1, b_s0 = lo - (r - 1);
                for (; y_acq < b_lo + b_n; ++y_acq) {
                    mbar_wait<RB_CTRL>(&yaempty[acq_slot], acq_ph, 54);
                    if (++acq_slot == G::RY) { acq_slot = 0; acq_ph ^= 1; }
                }
                const int slot_lo = (yslot0 + (lo - sg.y0)) % G::RY;
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t ad = adesc0 + (uint64_t)(smem_u32(s_x + xst * G::XROW) >> 4);
                    const RBPlan q = rb_plan<G::C, G::RY, BF16>(tmem_base, slot_lo, b_n, b_s0);
                    if (!SRB_DBG(p.dbg_skip & 2)) {
#pragma unroll
                        for (int kk = 0; kk < G::KS; ++kk) {
                            const int dx = kk >> 2, c2 = kk & 3;
                            rb_issue(q, ad + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4), bdesc0 + (uint64_t)((kk * G::W_STEP) >> 4));
                        }
                    }
                    umma_commit(&xempty[xst]);
                    // output row r-1 is complete (all three input rows seen); at the last x row also row r itself
                    if (r - 1 >= sg.y0) umma_commit(&yafull[(yslot0 + (r - 1 - sg.y0)) % G::RY]);
                    if (r == sg.hhi && r <= sg.y1 - 1) umma_commit(&yafull[(yslot0 + (r - sg.y0)) % G::RY]);
                }
                __syncwarp();
                if (++xst == G::XSTAGES) { xst = 0; xph ^= 1; }
            }
            const int ny = sg.y1 - sg.y0;
            ybase += ny;
            yslot0 = (yslot0 + ny) % G::RY;
        }
    } else if (warp >= 4) {
        // ===================== epilogue: two sets of 8 warps alternate rows; (+ LeakyReLU) + residual -> y =====================
        // A set = 4 lane quarters x 2 channel halves.  Residual in / result out straight between registers and global
        // memory: a warp's 32 lanes are 32 consecutive pixels = 512 contiguous bytes of each plane.
        const int lg = warp & 3;
        const int half = ((warp - 4) >> 2) & 1;    // channels 32*half .. : planes 4*half .. 4*half+3
        const int set = (warp - 4) >> 3;
        const int L = lg * 32 + lane;
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + half * 32 + ((uint32_t)(lg * 32) << 16);
        const bool has_res = p.res != nullptr && !SRB_DBG(p.dbg_skip & 4), no_st = SRB_DBG(p.dbg_skip & 4) != 0;
        pdl_wait();
        int c = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg)) {
            const RBSlot sl = rc_slot(p, sg.ss, grp);
            const int n = sl.n, gx = sl.gx;
            const bool valid = sl.out && gx * 8 + i < p.W;
            const size_t px0 = valid ? ((size_t)n * 8 + 4 * half) * HW + (size_t)(gx * 8 + i) : 0;
            const uint4* rg = reinterpret_cast<const uint4*>(p.res) + px0;
            uint4* yg = reinterpret_cast<uint4*>(p.y) + px0;
#pragma unroll 1
            for (int y = sg.y0; y < sg.y1; ++y, ++c) {
                if ((c & 1) != set) continue;
                const int slot = c % G::RY;
                // the residual is requested before the wait for the accumulator (a set has every other row: its wait is long)
                uint4 rq[4];
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    rq[j] = has_res && valid ? __ldg(rg + j * HW + (size_t)y * p.W) : make_uint4(0, 0, 0, 0);
                mbar_wait<RB_SLEEP>(&yafull[slot], (uint32_t)((c / G::RY) & 1), 55);
                tc_fence_after();
                uint32_t rr[32];
                tmem_ld32(t_lane + slot * G::C, rr);
                tmem_ld_wait();
                tmem_zero32(t_lane + slot * G::C);
                // (the bias index must be a compile-time constant to become a constant-bank operand: one copy per channel half)
                auto finish = [&](auto off) {
                    constexpr int OFF = decltype(off)::value;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t qq[4] = {rq[j].x, rq[j].y, rq[j].z, rq[j].w};
                        uint32_t ov[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpack2<BF16>(qq[k]);
                            float a = __uint_as_float(rr[8 * j + 2 * k]) + p.bias[OFF + 8 * j + 2 * k];
                            float b = __uint_as_float(rr[8 * j + 2 * k + 1]) + p.bias[OFF + 8 * j + 2 * k + 1];
                            if (p.lrelu) { a = fmaxf(a, a * p.slope); b = fmaxf(b, b * p.slope); }
                            ov[k] = pack2<BF16>(a + f.x, b + f.y);
                        }
                        if (valid && !no_st) yg[j * HW + (size_t)y * p.W] = make_uint4(ov[0], ov[1], ov[2], ov[3]);
                    }
                };
                if (half == 0) finish(std::integral_constant<int, 0>{});
                else finish(std::integral_constant<int, 32>{});
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&yaempty[slot]);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace srb
