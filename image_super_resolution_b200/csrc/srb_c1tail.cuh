// conv1 (3x3 64 -> 64, + bias, + long skip) and conv2 (9x9 64 -> 3, + bias, tanh -> uint8 / fp32) of the model.pt graph in ONE
// launch (SURVEY §8 a4 + a5 + a6; utils/models.py:690, :704, :691, :448-451): the row-fold conv1 kernel (srb_rowconv.cuh) and
// the row-fold tail kernel (srb_tail.cuh) composed through shared memory, so that conv1's 64-channel result — 128 B per pixel
// written and read back, 268 MB at the bench shape — never goes to HBM.  As two launches conv1 is bound by its three HBM
// streams (x, the skip, y: 402 MB = 62 us at the copy rate) and the tail re-reads what conv1 has just written.
//
// A CTA marches down its run of output rows.  Per row of 128 pixels (16 slots of 8 px, rb_slot with the tail's geometry):
//   x row (TMA, 10-px group boxes) -> conv1 MMAs, N = 3 x 64 over three adjacent row accumulators (ring of 3 x 64 TMEM columns)
//   -> mid epilogue: + bias + skip, rounded to the 16-bit operand type exactly as srb_rowconv stores it, written as the tail's
//      A operand [slot][plane][px] into a shared-memory ring (generic proxy + fence.proxy.async; no halo: the tail folds dx into N)
//   -> tail MMAs, N = 9 x 32 over nine adjacent row accumulators (ring of 10 x 32 TMEM columns: 192 + 320 = all 512)
//   -> tail epilogue: dx gather through shared memory, + bias, tanh, store.
// The 9x9 needs conv1 rows y0-4 .. y1+3 of a run of output rows [y0, y1): eight conv1 rows per CTA are computed twice (the
// price of not exchanging them between CTAs); conv1's ring is only three rows deep, so the MMAs of consecutive x rows are
// separated by the drain of the oldest row — the tail's MMAs (another warp, other accumulators) fill that gap.
// Results are bit-identical to srb_rowconv_fwd followed by srb_tail9_fwd (same K order per accumulator, same roundings).
#pragma once
#include "srb_rowconv.cuh"
#include "srb_tail.cuh"

namespace srb {

struct C1TailParams {
    int N, H, W, GX;
    int TR, TS, OV;           // tile geometry (rb_slot): the tail's (OV = 1)
    int total_rows;
    const uint8_t* w1;        // conv1 filters, packed as srb_rowconv's
    const uint8_t* w2;        // conv2 filters, packed as srb_tail9's
    float bias1[64];
    float bias2[3];
    const void* res;          // the long skip (conv0's output), C8 planes [N][8][H][W][8]
    void* y;                  // NCHW / NHWC uint8 or fp32 NCHW
    int flags;                // F_TANH, F_U8OUT (+ F_HWC_BGR), F_F32OUT
};

#ifndef CT_MID_WAIT
#define CT_MID_WAIT RB_SLEEP_NS      // the mid epilogue's wait for a conv1 row: ns between polls (< 0: parked by the hardware)
#endif
struct CTGeom {
    static constexpr int C = 64;
    static constexpr int XBOX = 8 * 10 * 16, XROW = 16 * XBOX, XSTAGES = 3;      // x rows with their halo: 20 KB each
    static constexpr int MBOX = 8 * 8 * 16, MROW = 16 * MBOX, MSLOTS = 2;        // conv1 rows as the tail's operand: 16 KB each
    static constexpr int KS1 = RCGeom::KS, W1_STEP = RCGeom::W_STEP, W1_BYTES = RCGeom::W_BYTES;
    static constexpr int KS2 = TLGeom::KS, W2_STEP = TLGeom::W_STEP, W2_BYTES = TLGeom::W_BYTES, NS = TLGeom::NS;
    static constexpr int RY1 = 3, RY2 = 10;                    // accumulator rings: 3 x 64 + 10 x 32 = 512 TMEM columns
    static constexpr int TM_T = RY1 * C;                       // first column of the tail ring
    static constexpr int PLEN = 144, PCOLS = 27;               // exchange buffer (rows of >= 64 px: two image rows + pads fit 144)
    static constexpr int P_BYTES = PCOLS * PLEN * 4;
    static constexpr int OFF_W1 = 0, OFF_W2 = W1_BYTES, OFF_X = OFF_W2 + W2_BYTES, OFF_M = OFF_X + XSTAGES * XROW;
    static constexpr int OFF_P = OFF_M + MSLOTS * MROW, OFF_BAR = OFF_P + P_BYTES;
    static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 512;
    static constexpr int THREADS = 512;                        // 4 control warps, 8 mid-epilogue warps, 4 tail-epilogue warps
    static_assert(RY1 * C + RY2 * NS == 512, "the accumulator rings fill TMEM");
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
};

template <bool BF16>
__global__ void __launch_bounds__(CTGeom::THREADS, 1) c1tail_kernel(const __grid_constant__ CUtensorMap tmap_x, const C1TailParams p) {
    using G = CTGeom;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* s_w1 = smem + G::OFF_W1;
    uint8_t* s_w2 = smem + G::OFF_W2;
    uint8_t* s_x = smem + G::OFF_X;
    uint8_t* s_m = smem + G::OFF_M;
    float* const P = reinterpret_cast<float*>(smem + G::OFF_P);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + G::OFF_BAR);
    uint64_t* xfull = bars;             // [XSTAGES] TMA -> conv1 issuer
    uint64_t* xempty = bars + 4;        // [XSTAGES] conv1 issuer -> TMA
    uint64_t* y1full = bars + 8;        // [RY1] conv1 row complete            conv1 issuer -> mid epilogue
    uint64_t* y1empty = bars + 12;      // [RY1] drained and re-armed          mid epilogue -> conv1 issuer
    uint64_t* mfull = bars + 16;        // [MSLOTS] conv1 row in shared memory mid epilogue -> tail issuer
    uint64_t* mempty = bars + 18;       // [MSLOTS]                            tail issuer -> mid epilogue
    uint64_t* y2full = bars + 20;       // [RY2] output row complete           tail issuer -> tail epilogue
    uint64_t* y2empty = bars + 30;      // [RY2] drained and re-armed          tail epilogue -> tail issuer
    uint64_t* wbar = bars + 40;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 41);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < G::PCOLS * G::PLEN; i += G::THREADS) P[i] = 0.f;        // pads between image rows stay zero
    if (threadIdx.x == 0) {
        for (int i = 0; i < G::XSTAGES; ++i) { mbar_init(&xfull[i], 1); mbar_init(&xempty[i], 1); }
        for (int i = 0; i < G::RY1; ++i) { mbar_init(&y1full[i], 1); mbar_init(&y1empty[i], 8); }
        for (int i = 0; i < G::MSLOTS; ++i) { mbar_init(&mfull[i], 8); mbar_init(&mempty[i], 1); }
        for (int i = 0; i < G::RY2; ++i) { mbar_init(&y2full[i], 1); mbar_init(&y2empty[i], 4); }
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
    if (warp >= 4 && warp < 8) {       // every MMA accumulates: all rows start at zero (the epilogues add the biases)
        const uint32_t lgz = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll 1
        for (int c = 0; c < 512; c += 32) tmem_zero32(tmem_base + lgz + c);
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    pdl_launch_dependents();

    const int g_begin = (int)((long long)blockIdx.x * p.total_rows / gridDim.x);
    const int g_end = (int)((long long)(blockIdx.x + 1) * p.total_rows / gridDim.x);
    const size_t HW = (size_t)p.H * p.W;
    // a segment (TLSeg): output rows [y0, y1) of strip-set ss; conv1 rows xlo .. xhi (= y0-4 .. y1+3, clipped); x rows one more each way
    auto x_lo = [&](const TLSeg& s) { return s.xlo > 0 ? s.xlo - 1 : 0; };
    auto x_hi = [&](const TLSeg& s) { return s.xhi + 1 < p.H ? s.xhi + 1 : p.H - 1; };

    if (warp == 0) {
        // ===================== TMA producer: both filters once, then one x row (16 group boxes with halo) per stage =====================
        if (elect_one()) {
            mbar_arrive_expect_tx(wbar, G::W1_BYTES + G::W2_BYTES);
            for (int off = 0; off < G::W1_BYTES; off += 12288) bulk_load_1d(s_w1 + off, p.w1 + off, 12288, wbar);
            for (int off = 0; off < G::W2_BYTES; off += 12288) bulk_load_1d(s_w2 + off, p.w2 + off, 12288, wbar);
        }
        __syncwarp();
        pdl_wait();
        int xc = 0, g = g_begin;
        TLSeg sg;
        while (tl_next_seg(p, g, g_end, sg)) {
            const RBSlot sl = rb_slot(p, sg.ss, lane & 15);
            const int n = sl.n, gx = sl.gx;
            const int xl = x_lo(sg), xh = x_hi(sg);
#pragma unroll 1
            for (int r = xl; r <= xh; ++r, ++xc) {
                const int st = xc % G::XSTAGES;
                mbar_wait<RB_CTRL>(&xempty[st], (uint32_t)(((xc / G::XSTAGES) & 1) ^ 1), 81);
                if (lane == 0) mbar_arrive_expect_tx(&xfull[st], G::XROW);
                __syncwarp();
                if (lane < 16)
                    tma_load_4d(s_x + st * G::XROW + lane * G::XBOX, &tmap_x, &xfull[st], (gx * 8 - 1) * 4, r, 0, n);
                __syncwarp();
            }
        }
    } else if (warp == 1) {
        // ===================== conv1 MMA issuer: one x row per step, N = 3 x 64 into the three-row ring =====================
        const uint64_t adesc0 = umma_desc_nosw(0, 160, G::XBOX);
        const uint64_t bdesc0 = umma_desc_nosw(smem_u32(s_w1), 3 * G::C * 16, 128);
        mbar_wait(wbar, 0, 82);
        int xst = 0; uint32_t xph = 0;
        int yslot0 = 0;                                // ring slot of the segment's first conv1 row
        int acq_slot = 0; uint32_t acq_ph = 1;
        int ybase = 0, y_acq = 0, g = g_begin;
        TLSeg sg;
        while (tl_next_seg(p, g, g_end, sg)) {
            const int m0 = sg.xlo, m1 = sg.xhi + 1;    // conv1 rows [m0, m1)
            const int xl = x_lo(sg), xh = x_hi(sg);
#pragma unroll 1
            for (int r = xl; r <= xh; ++r) {
                mbar_wait<RB_CTRL>(&xfull[xst], xph, 83);
                const int lo = r - 1 > m0 ? r - 1 : m0, hi = r + 1 < m1 - 1 ? r + 1 : m1 - 1;
                const int b_lo = ybase + lo - m0, b_n = hi - lo + 1, b_s0 = lo - (r - 1);
                for (; y_acq < b_lo + b_n; ++y_acq) {
                    mbar_wait<RB_CTRL>(&y1empty[acq_slot], acq_ph, 84);
                    if (++acq_slot == G::RY1) { acq_slot = 0; acq_ph ^= 1; }
                }
                const int slot_lo = (yslot0 + (lo - m0)) % G::RY1;
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t ad = adesc0 + (uint64_t)(smem_u32(s_x + xst * G::XROW) >> 4);
                    const RBPlan q = rb_plan<G::C, G::RY1, BF16>(tmem_base, slot_lo, b_n, b_s0);
#pragma unroll
                    for (int kk = 0; kk < G::KS1; ++kk) {
                        const int dx = kk >> 2, c2 = kk & 3;
                        rb_issue(q, ad + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4), bdesc0 + (uint64_t)((kk * G::W1_STEP) >> 4));
                    }
                    umma_commit(&xempty[xst]);
                    // conv1 row r-1 is complete (all three x rows seen); at the last x row also row r itself
                    if (r - 1 >= m0) umma_commit(&y1full[(yslot0 + (r - 1 - m0)) % G::RY1]);
                    if (r == xh && r <= m1 - 1) umma_commit(&y1full[(yslot0 + (r - m0)) % G::RY1]);
                }
                __syncwarp();
                if (++xst == G::XSTAGES) { xst = 0; xph ^= 1; }
            }
            const int nm = m1 - m0;
            ybase += nm;
            yslot0 = (yslot0 + nm) % G::RY1;
        }
    } else if (warp == 2) {
        // ===================== tail MMA issuer: one conv1 row per step, N = 9 x 32 (two MMAs) into the ten-row ring =====================
        const uint64_t adesc0 = umma_desc_nosw(0, 128, G::MBOX);            // planes 128 B apart, groups MBOX apart
        const uint64_t bdesc0 = umma_desc_nosw(smem_u32(s_w2), TLGeom::BROWS * 16, 128);
        const uint32_t tm_t = tmem_base + G::TM_T;
        mbar_wait(wbar, 0, 85);
        int mc = 0;                                    // running conv1-row counter (shared-memory ring)
        int yslot0 = 0;
        int acq_slot = 0; uint32_t acq_ph = 1;
        int ybase = 0, y_acq = 0, g = g_begin;
        TLSeg sg;
        while (tl_next_seg(p, g, g_end, sg)) {
#pragma unroll 1
            for (int r = sg.xlo; r <= sg.xhi; ++r, ++mc) {
                const int ms = mc % G::MSLOTS;
                mbar_wait<RB_CTRL>(&mfull[ms], (uint32_t)((mc / G::MSLOTS) & 1), 86);
                const int lo = r - 4 > sg.y0 ? r - 4 : sg.y0, hi = r + 4 < sg.y1 - 1 ? r + 4 : sg.y1 - 1;
                const int n = hi - lo + 1, s0 = lo - (r - 4);              // slice s <-> output row r-4+s <-> dy = 8-s
                for (; y_acq < ybase + hi - sg.y0 + 1; ++y_acq) {
                    mbar_wait<RB_CTRL>(&y2empty[acq_slot], acq_ph, 87);
                    if (++acq_slot == G::RY2) { acq_slot = 0; acq_ph ^= 1; }
                }
                const int slot_lo = (yslot0 + (lo - sg.y0)) % G::RY2;
                tc_fence_after();
                if (elect_one()) {
                    const uint64_t ad = adesc0 + (uint64_t)(smem_u32(s_m + ms * G::MROW) >> 4);
                    // at most two pieces: N <= 256 (8 rows) and the ring's wrap-around (a piece cut at 8 rows leaves one row, which
                    // fits before the wrap: 9 rows in a ring of 10)
                    int n1 = n < 8 ? n : 8;
                    if (n1 > G::RY2 - slot_lo) n1 = G::RY2 - slot_lo;
                    const int n2 = n - n1, slot2 = (slot_lo + n1) % G::RY2;
                    const uint32_t d1 = tm_t + slot_lo * G::NS, d2 = tm_t + slot2 * G::NS;
                    const uint32_t id1 = umma_idesc_f16(128, n1 * G::NS, BF16 ? 1 : 0);
                    const uint32_t id2 = umma_idesc_f16(128, (n2 > 0 ? n2 : 1) * G::NS, BF16 ? 1 : 0);
                    const uint64_t b1 = bdesc0 + (uint64_t)((s0 * G::NS * 16) >> 4), b2 = bdesc0 + (uint64_t)(((s0 + n1) * G::NS * 16) >> 4);
#pragma unroll
                    for (int kk = 0; kk < G::KS2; ++kk) {
                        const uint64_t a = ad + (uint64_t)((2 * kk * 128) >> 4), bo = (uint64_t)((kk * G::W2_STEP) >> 4);
                        umma_f16_ss(d1, a, b1 + bo, id1, 1u);
                        if (n2 > 0) umma_f16_ss(d2, a, b2 + bo, id2, 1u);
                    }
                    umma_commit(&mempty[ms]);
                    // output row r-4 has seen its last conv1 row; at the segment's last conv1 row so have all that remain
                    if (r - 4 >= sg.y0) umma_commit(&y2full[(yslot0 + (r - 4 - sg.y0)) % G::RY2]);
                    if (r == sg.xhi)
                        for (int yo = (r - 3 > sg.y0 ? r - 3 : sg.y0); yo < sg.y1; ++yo) umma_commit(&y2full[(yslot0 + (yo - sg.y0)) % G::RY2]);
                }
                __syncwarp();
            }
            const int ny = sg.y1 - sg.y0;
            ybase += ny;
            yslot0 = (yslot0 + ny) % G::RY2;
        }
    } else if (warp >= 4 && warp < 12) {
        // ===================== mid epilogue (8 warps, one conv1 row at a time): + bias + skip -> 16-bit -> the tail's operand =====================
        const int lg = warp & 3;
        const int half = ((warp - 4) >> 2) & 1;    // channels 32*half .. : planes 4*half .. 4*half+3
        const int L = lg * 32 + lane;
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + half * 32 + ((uint32_t)(lg * 32) << 16);
        const uint32_t m_lane = smem_u32(s_m) + (uint32_t)(grp * G::MBOX + 4 * half * 128 + i * 16);
        pdl_wait();
        int c = 0, g = g_begin;
        TLSeg sg;
        while (tl_next_seg(p, g, g_end, sg)) {
            const RBSlot sl = rb_slot(p, sg.ss, grp);
            const bool valid = sl.in && sl.gx * 8 + i < p.W;
            const size_t px0 = valid ? ((size_t)sl.n * 8 + 4 * half) * HW + (size_t)(sl.gx * 8 + i) : 0;
            const uint4* rg = reinterpret_cast<const uint4*>(p.res) + px0;
            // (one set of warps takes every row: the skip of row m + 1 is requested while row m is finished, or its DRAM round
            // trip would sit between the drain and the shared-memory stores of every row; two rows of lead measured slower:
            // 108.2 against 104.0 us)
            uint4 nxt[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) nxt[j] = valid ? ldg_nc16(rg + j * HW + (size_t)sg.xlo * p.W) : make_uint4(0, 0, 0, 0);
#pragma unroll 1
            for (int m = sg.xlo; m <= sg.xhi; ++m, ++c) {
                const int slot = c % G::RY1, ms = c % G::MSLOTS;
                uint4 rq[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) rq[j] = nxt[j];
                mbar_wait<CT_MID_WAIT>(&y1full[slot], (uint32_t)((c / G::RY1) & 1), 88);
                tc_fence_after();
                uint32_t rr[32];
                tmem_ld32(t_lane + slot * G::C, rr);
                tmem_ld_wait();
                tmem_zero32(t_lane + slot * G::C);
                tmem_st_wait();                    // the accumulator row goes back at once: conv1's ring has no spare row
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&y1empty[slot]);
                if (m < sg.xhi && valid) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) nxt[j] = ldg_nc16(rg + j * HW + (size_t)(m + 1) * p.W);
                }
                uint4 o[4];
                auto finish = [&](auto off) {
                    constexpr int OFF = decltype(off)::value;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t qq[4] = {rq[j].x, rq[j].y, rq[j].z, rq[j].w};
                        uint32_t ov[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = unpack2<BF16>(qq[k]);
                            const float a = __uint_as_float(rr[8 * j + 2 * k]) + p.bias1[OFF + 8 * j + 2 * k];
                            const float b = __uint_as_float(rr[8 * j + 2 * k + 1]) + p.bias1[OFF + 8 * j + 2 * k + 1];
                            ov[k] = pack2<BF16>(a + f.x, b + f.y);
                        }
                        o[j] = make_uint4(ov[0], ov[1], ov[2], ov[3]);
                    }
                };
                if (half == 0) finish(std::integral_constant<int, 0>{});
                else finish(std::integral_constant<int, 32>{});
                mbar_wait<RB_SLEEP>(&mempty[ms], (uint32_t)(((c / G::MSLOTS) & 1) ^ 1), 89);
#pragma unroll
                for (int j = 0; j < 4; ++j) sts_u4(m_lane + ms * G::MROW + j * 128, o[j]);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&mfull[ms]);
            }
        }
    } else if (warp >= 12) {
        // ===================== tail epilogue (4 warps): dx gather through shared memory, + bias, tanh, store =====================
        const int lg = warp & 3;
        const int L = lg * 32 + lane;
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + G::TM_T + ((uint32_t)(lg * 32) << 16);
        const int gxw = p.TS > 0 ? 128 : p.GX * 8;                  // pixels of one image row in the tile
        const int pidx = (L / gxw) * (gxw + 8) + (L % gxw);         // padded position of this lane's pixel (its own column at +4)
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
                const int slot = c % G::RY2;
                mbar_wait<RB_SLEEP>(&y2full[slot], (uint32_t)((c / G::RY2) & 1), 90);
                tc_fence_after();
                uint32_t rr[32];
                tmem_ld32(t_lane + slot * G::NS, rr);
                tmem_ld_wait();
                tmem_zero32(t_lane + slot * G::NS);
#pragma unroll
                for (int k = 0; k < G::PCOLS; ++k) P[k * G::PLEN + pidx + 4] = inside ? __uint_as_float(rr[k]) : 0.f;
                tmem_st_wait();
                tc_fence_before();
                named_bar_sync(1, 128);
                if (lane == 0) mbar_arrive(&y2empty[slot]);
                float acc[3] = {p.bias2[0], p.bias2[1], p.bias2[2]};
#pragma unroll
                for (int dx = 0; dx < 9; ++dx)
#pragma unroll
                    for (int co = 0; co < 3; ++co) acc[co] += P[(dx * 3 + co) * G::PLEN + pidx + dx];
                named_bar_sync(1, 128);           // the next row overwrites P
                if (valid) {
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
