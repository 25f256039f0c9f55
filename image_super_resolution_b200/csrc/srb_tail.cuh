// Row-fold 9x9 conv 64 -> 3 (+ bias, tanh, uint8 / fp32 NCHW output) on tcgen05: conv2 + TanhToArrayImage of the
// model.pt graph (utils/models.py:691, :448-451).
//
// With Cout = 3 the conv has almost no N to offer the tensor core, so BOTH tap directions are folded into N:
//   * dx: column dx*3+co of an accumulator row holds the contribution of input pixel x' to output pixel x'-dx+4
//     (the epilogue gathers the 9 shifted partial sums — through shared memory, they cross warps);
//   * dy: the MMA's D operand spans the 9 adjacent row accumulators of output rows r-4 .. r+4 of input row r
//     (B = [W(dy=8) | ... | W(dy=0)], 9 x 32 columns; issued as two MMAs, N <= 256).
// Per 128 input pixels and 16 input channels that is 2 MMAs (N = 256 + 32) instead of the tiled kernel's 18 MMAs of
// N = 32 on a tile whose half is halo: 4 K-steps per row, ~600 clk per 128 px against 5 800.  The M-tile needs no halo at
// all: the A operand is 16 plain group boxes [8 planes][8 px] of one row, out-of-image taps are simply never added.
// A CTA marches down a run of rows; accumulator rows live in a ring of 16 x 32 TMEM columns.
#pragma once
#include "srb_resblock.cuh"

namespace srb {

struct TailParams {
    int N, H, W, GX;
    int TR, TS, OV;           // tile geometry (rb_slot); OV = 1: the dx exchange needs the neighbouring groups' partial sums
    int total_rows;
    const uint8_t* w;         // packed [4 K-steps][2 granules][288 rows = 9 dy-slices x 32][8 k] 16-bit
    float bias[3];
    void* y;                  // NCHW uint8 (F_U8OUT) or fp32 (F_F32OUT)
    int flags;                // F_TANH, F_U8OUT (+ F_HWC_BGR), F_F32OUT
    int row5d;                // the tensor map is 5-D [32 words][8 planes][groups][rows][images]: ONE box per input row (W % 8 == 0)
    int dbg_skip;
};

#ifndef TL_XSTAGES
#define TL_XSTAGES 6
#endif
#ifndef TL_SETS
#define TL_SETS 2
#endif
struct TLGeom {
    static constexpr int C = 64, NS = 32, TAPS = 9;            // input channels, columns per row accumulator, dy taps
    static constexpr int XBOX = 8 * 8 * 16, XROW = 16 * XBOX, XSTAGES = TL_XSTAGES;       // 1 KB boxes, 16 KB per row
    static constexpr int KS = C / 16;                          // 4 K-steps per row
    static constexpr int BROWS = TAPS * NS;                    // 288 B rows per granule
    static constexpr int W_STEP = 2 * BROWS * 16;              // 9216
    static constexpr int W_BYTES = KS * W_STEP;                // 36864
    static constexpr int RY = 16;                              // accumulator ring: 16 x 32 = 512 TMEM columns
    static constexpr int PLEN = 256, PCOLS = 27;               // exchange buffer [27 columns][256 padded pixels] fp32 per set
    static constexpr int P_BYTES = PCOLS * PLEN * 4;           // 27 648
    static constexpr int SETS = TL_SETS;                        // epilogue sets of 4 warps alternating output rows
    static constexpr int OFF_W = 0, OFF_X = W_BYTES, OFF_P = OFF_X + XSTAGES * XROW, OFF_BAR = OFF_P + SETS * P_BYTES;
    static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 512;
    static constexpr int THREADS = 128 + 128 * SETS;            // 4 control warps + SETS epilogue sets of 4 warps
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

// output rows [y0, y1) of strip-set ss and the input rows they need
struct TLSeg { int ss, y0, y1, xlo, xhi; };
template <class Params>
__device__ __forceinline__ bool tl_next_seg(const Params& p, int& g, int g1, TLSeg& s) {
    if (g >= g1) return false;
    s.ss = g / p.H;
    s.y0 = g - s.ss * p.H;
    const int n = g1 - g;
    s.y1 = s.y0 + n < p.H ? s.y0 + n : p.H;
    s.xlo = s.y0 - 4 > 0 ? s.y0 - 4 : 0;
    s.xhi = s.y1 + 3 < p.H - 1 ? s.y1 + 3 : p.H - 1;
    g += s.y1 - s.y0;
    return true;
}

template <bool BF16>
__global__ void __launch_bounds__(TLGeom::THREADS, 1) tail9_kernel(const __grid_constant__ CUtensorMap tmap_x, const TailParams p) {
    using G = TLGeom;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);     // keeps the shared address space
    uint8_t* s_w = smem + G::OFF_W;
    uint8_t* s_x = smem + G::OFF_X;
    float* s_p = reinterpret_cast<float*>(smem + G::OFF_P);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
    uint64_t* xfull = bars;            // [XSTAGES] TMA -> MMA
    uint64_t* xempty = bars + 8;       // [XSTAGES] MMA -> TMA
    uint64_t* yafull = bars + 16;      // [RY] accumulator row complete      MMA -> epilogue
    uint64_t* yaempty = bars + 32;     // [RY] drained and zeroed            epilogue -> MMA
    uint64_t* wbar = bars + 48;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 49);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < G::SETS * G::PCOLS * G::PLEN; i += G::THREADS) s_p[i] = 0.f;   // pads between image rows stay zero
    if (p.row5d && p.TS == 0 && p.GX * p.TR < 16)      // slots the row box never fills: their operand rows read as zeros
        for (int i = threadIdx.x; i < G::XSTAGES * G::XROW / 16; i += G::THREADS) reinterpret_cast<uint4*>(s_x)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) {
        for (int i = 0; i < G::XSTAGES; ++i) { mbar_init(&xfull[i], 1); mbar_init(&xempty[i], 1); }
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
    if (warp >= 4 && warp < 8) {       // every MMA accumulates: all rows start at zero
        const uint32_t lg = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll 1
        for (int c = 0; c < 512; c += 32) tmem_zero32(tmem_base + lg + c);
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
        // ===================== TMA producer: filters once, then one input row (16 group boxes, no halo) per stage =====================
        if (elect_one()) {
            mbar_arrive_expect_tx(wbar, G::W_BYTES);
            for (int off = 0; off < G::W_BYTES; off += 12288) bulk_load_1d(s_w + off, p.w + off, 12288, wbar);
        }
        __syncwarp();
        pdl_wait();
        int xc = 0, g = g_begin;
        const uint32_t xbytes5 = p.TS == 0 ? (uint32_t)(p.TR * p.GX * G::XBOX) : (uint32_t)G::XROW;
        TLSeg sg;
        while (tl_next_seg(p, g, g_end, sg)) {
            const RBSlot sl = rb_slot(p, sg.ss, lane & 15);
            const RBSlot sl0 = rb_slot(p, sg.ss, 0);
            const int n = sl.n, gx = sl.gx;
#pragma unroll 1
            for (int r = sg.xlo; r <= sg.xhi; ++r, ++xc) {
                const int st = xc % G::XSTAGES;
                mbar_wait<RB_CTRL>(&xempty[st], (uint32_t)(((xc / G::XSTAGES) & 1) ^ 1), 61);
                if SRB_DBG(p.dbg_skip & 1) { if (lane == 0) mbar_arrive(&xfull[st]); __syncwarp(); continue; }
                if (p.row5d) {
                    // the tile's slots are consecutive groups of TR image rows (narrow) or of one strip (wide): one 5-D box
                    // [8 px][8 planes][groups][1 row][images] lands as [slot][plane][px] — 1 TMA instruction per row instead of 16
                    // (the producer warp pays ~60-90 clk per instruction: 16 of them set the row period of this kernel)
                    if (lane == 0) {
                        mbar_arrive_expect_tx(&xfull[st], xbytes5);
                        tma_load_5d(s_x + st * G::XROW, &tmap_x, &xfull[st], 0, 0, sl0.gx, r, sl0.n);
                    }
                    __syncwarp();
                    continue;
                }
                if (lane == 0) mbar_arrive_expect_tx(&xfull[st], G::XROW);
                __syncwarp();
                if (lane < 16)
                    tma_load_4d(s_x + st * G::XROW + lane * G::XBOX, &tmap_x, &xfull[st], gx * 8 * 4, r, 0, n);
                __syncwarp();
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: one input row per step, N = 9 x 32 (two MMAs) into the accumulator ring =====================
        const uint64_t adesc0 = umma_desc_nosw(0, 128, G::XBOX);            // planes 128 B apart, groups XBOX apart
        const uint64_t bdesc0 = umma_desc_nosw(smem_u32(s_w), G::BROWS * 16, 128);
        mbar_wait(wbar, 0, 62);
        int xst = 0; uint32_t xph = 0;
        int yslot0 = 0;                                // ring slot of the segment's first output row
        int acq_slot = 0; uint32_t acq_ph = 1;
        int ybase = 0, y_acq = 0, g = g_begin;
        TLSeg sg;
        while (tl_next_seg(p, g, g_end, sg)) {
#pragma unroll 1
            for (int r = sg.xlo; r <= sg.xhi; ++r) {
                mbar_wait<RB_CTRL>(&xfull[xst], xph, 63);
                const int lo = r - 4 > sg.y0 ? r - 4 : sg.y0, hi = r + 4 < sg.y1 - 1 ? r + 4 : sg.y1 - 1;
                const int n = hi - lo + 1, s0 = lo - (r - 4);              // slice s <-> output row r-4+s <-> dy = 8-s
                for (; y_acq < ybase + hi - sg.y0 + 1; ++y_acq) {
                    mbar_wait<RB_CTRL>(&yaempty[acq_slot], acq_ph, 64);
                    if (++acq_slot == G::RY) { acq_slot = 0; acq_ph ^= 1; }
                }
                const int slot_lo = (yslot0 + (lo - sg.y0)) % G::RY;
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t ad = adesc0 + (uint64_t)(smem_u32(s_x + xst * G::XROW) >> 4);
                    // at most two pieces: N <= 256 (8 rows) and the ring's wrap-around
                    int n1 = n < 8 ? n : 8;
                    if (n1 > G::RY - slot_lo) n1 = G::RY - slot_lo;
                    const int n2 = n - n1, slot2 = (slot_lo + n1) % G::RY;
                    const uint32_t d1 = tmem_base + slot_lo * G::NS, d2 = tmem_base + slot2 * G::NS;
                    const uint32_t id1 = umma_idesc_f16(128, n1 * G::NS, BF16 ? 1 : 0);
                    const uint32_t id2 = umma_idesc_f16(128, (n2 > 0 ? n2 : 1) * G::NS, BF16 ? 1 : 0);
                    const uint64_t b1 = bdesc0 + (uint64_t)((s0 * G::NS * 16) >> 4), b2 = bdesc0 + (uint64_t)(((s0 + n1) * G::NS * 16) >> 4);
                    if (!SRB_DBG(p.dbg_skip & 2)) {
#pragma unroll
                        for (int kk = 0; kk < G::KS; ++kk) {
                            const uint64_t a = ad + (uint64_t)((2 * kk * 128) >> 4), bo = (uint64_t)((kk * G::W_STEP) >> 4);
                            umma_f16_ss(d1, a, b1 + bo, id1, 1u);
                            if (n2 > 0) umma_f16_ss(d2, a, b2 + bo, id2, 1u);
                        }
                    }
                    umma_commit(&xempty[xst]);
                    // output row r-4 has seen its last input row; at the segment's last input row so have all that remain
                    if (r - 4 >= sg.y0) umma_commit(&yafull[(yslot0 + (r - 4 - sg.y0)) % G::RY]);
                    if (r == sg.xhi)
                        for (int yo = (r - 3 > sg.y0 ? r - 3 : sg.y0); yo < sg.y1; ++yo) umma_commit(&yafull[(yslot0 + (yo - sg.y0)) % G::RY]);
                }
                __syncwarp();
                if (++xst == G::XSTAGES) { xst = 0; xph ^= 1; }
            }
            const int ny = sg.y1 - sg.y0;
            ybase += ny;
            yslot0 = (yslot0 + ny) % G::RY;
        }
    } else if (warp >= 4) {
        // ===================== epilogue: two sets of 4 warps alternate output rows =====================
        // accumulator row -> shared-memory exchange (column dx*3+co of input pixel x' belongs to output pixel x'-dx+4) ->
        // gather 9 x 3 partials -> + bias -> tanh -> uint8 / fp32 NCHW
        const int lg = warp & 3;
        const int set = (warp - 4) >> 2;
        const int L = lg * 32 + lane;
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + ((uint32_t)(lg * 32) << 16);
        float* const P = s_p + set * (G::PCOLS * G::PLEN);
        const int gxw = p.TS > 0 ? 128 : p.GX * 8;                  // pixels of one image row in the tile
        const int pidx = (L / gxw) * (gxw + 8) + (L % gxw);         // padded position of this lane's pixel (its own column at +4)
        const bool no_st = SRB_DBG(p.dbg_skip & 4) != 0;
        pdl_wait();
        int c = 0, g = g_begin;
        TLSeg sg;
        while (tl_next_seg(p, g, g_end, sg)) {
            const RBSlot sl = rb_slot(p, sg.ss, grp);
            const int n = sl.n, gx = sl.gx;
            const int x = gx * 8 + i;
            const bool inside = sl.in && x < p.W, valid = sl.out && x < p.W;
#pragma unroll 1
            for (int y = sg.y0; y < sg.y1; ++y, ++c) {
                if ((c % G::SETS) != set) continue;
                const int slot = c % G::RY;
                mbar_wait<RB_SLEEP>(&yafull[slot], (uint32_t)((c / G::RY) & 1), 65);
                tc_fence_after();
                uint32_t rr[32];
                tmem_ld32(t_lane + slot * G::NS, rr);
                tmem_ld_wait();
                tmem_zero32(t_lane + slot * G::NS);
#pragma unroll
                for (int k = 0; k < G::PCOLS; ++k) P[k * G::PLEN + pidx + 4] = inside ? __uint_as_float(rr[k]) : 0.f;
                tmem_st_wait();
                tc_fence_before();
                named_bar_sync(1 + set, 128);
                if (lane == 0) mbar_arrive(&yaempty[slot]);
                float acc[3] = {p.bias[0], p.bias[1], p.bias[2]};
#pragma unroll
                for (int dx = 0; dx < 9; ++dx)
#pragma unroll
                    for (int co = 0; co < 3; ++co) acc[co] += P[(dx * 3 + co) * G::PLEN + pidx + dx];
                named_bar_sync(1 + set, 128);           // the set's next row overwrites P
                if (valid && !no_st) {
#pragma unroll
                    for (int co = 0; co < 3; ++co) {
                        float tt = acc[co];
                        if (p.flags & F_TANH) tt = tanhf(tt);
                        const size_t o = (p.flags & F_HWC_BGR) ? ((size_t)n * HW + (size_t)y * p.W + x) * 3 + (2 - co)
                                                               : ((size_t)n * 3 + co) * HW + (size_t)y * p.W + x;
                        if (p.flags & F_U8OUT) {
                            float u = (tt + 1.0f) * 0.5f;
                            u = u * 255.0f;
                            u = fminf(fmaxf(rintf(u), 0.f), 255.f);
                            reinterpret_cast<uint8_t*>(p.y)[o] = (uint8_t)u;
                        } else {
                            reinterpret_cast<float*>(p.y)[o] = tt;
                        }
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace srb
