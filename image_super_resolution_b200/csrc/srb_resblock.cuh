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
// Warp roles (640 threads): 0 TMA producer, 1 convA MMA issuer, 2 convB MMA issuer, 4-11 hidden epilogue, 12-19 output
// epilogue (each as two sets of 4 warps that alternate rows).  The tensor pipe interleaves the two issuers' independent accumulation chains.
#pragma once
#include <type_traits>
#include <cuda.h>
#include "srb_ptx.cuh"
#include "srb_conv.cuh"

namespace srb {

struct ResblockParams {
    int N, H, W, GX;          // GX = groups (8 px) per image row
    int TR, TS, OV;           // tile geometry (rb_slot): image rows per tile (narrow), strips per image row (wide), overlap
    int total_rows;           // strip-sets x H: the flattened (strip-set, output row) space split evenly over CTAs
    const uint8_t* wA;        // fold-packed 64 -> 32:  12 K-steps x [2][96][8]
    const uint8_t* wB;        // fold-packed 32 -> 64:   6 K-steps x [2][192][8]
    const uint8_t* wA2;       // CTA-pair packing (pack_fold_pair): [rank][4 arrangements] of each CTA's half of the B rows
    const uint8_t* wB2;
    float bias[32 + 64];      // convA, convB biases, by value (copied to shared memory in the prologue)
    const void* x;            // [N][8][H][W][8]
    void* y;                  // [N][8][H][W][8]
    float slope;
    long long* dbg_trace;     // optional clock stamps of CTA 0: [32 rows][32]
    int dbg_skip;             // bottleneck probes: 1 = no TMA loads, 2 = no MMAs (16 = no convA, 32 = no convB MMAs), 4 = no global I/O in the output epilogue
};

// RB_KSPLIT: convA is issued by TWO warps, each owning half of the input channels (K) and its own hidden accumulator ring;
// the hidden epilogue adds the two partial sums.  Why (profiles/ubench_mix_r2.log): a chain of N = 96 MMAs into one
// accumulator runs at 73 clk per MMA (48 clk of math) whatever the number of accumulators one thread alternates between,
// while two issuers with a chain each reach 67 clk — and in the kernel the issuing thread's per-row bookkeeping is exposed
// because the tensor pipe's queue is only a few MMAs deep: with two issuers the pipe stays fed while either thread waits.
// The two rings double the hidden TMEM columns, so all rings are 4 rows deep (2*4*32 + 4*64 = 512 columns).
// Measured: fused block 89.1 -> 83.1 us per launch in the bench, forward 1180 -> 1210 Mpix/s.
#ifndef RB_KSPLIT
#define RB_KSPLIT 1
#endif
#ifndef RB_RH
#if RB_KSPLIT
#define RB_RH 4
#define RB_RY 4
#else
#define RB_RH 6
#define RB_RY 5
#endif
#endif
#ifndef RB_HSLOTS
#define RB_HSLOTS 2            // hidden rows in shared memory: 2 vs 3 slots measure equal (profiles/ab_hslots_r2.log); the third paid for an x stage
#endif
#ifndef RB_XSTAGES
#define RB_XSTAGES 4
#endif
#ifndef RB_XSTAGES2
#define RB_XSTAGES2 4          // x ring of the CTA-pair kernel (its filters take 120 KB): 3 -> 4 stages = forward 798.2 -> 787.8 us;
                               // ncu's samples had the convA issuers waiting 24 % of their time for an x row with three stages
#endif
#ifndef RB_BIAS_MODE
#define RB_BIAS_MODE 1       // 0: accumulator rows armed with the bias (re-armed from shared memory); 1: rows armed with zero,
#endif                       //    the epilogues add the bias as a constant-bank operand (no shared-memory reads, no registers)
#ifndef RB_SLEEP_NS
#define RB_SLEEP_NS 40
#endif
#ifndef RB_CTRL_NS
#define RB_CTRL_NS 0
#endif
constexpr int RB_SLEEP = RB_SLEEP_NS;   // epilogue warps' waits: > 0 ns between polls, < 0 parked by the hardware up to that long
constexpr int RB_CTRL = RB_CTRL_NS;     // producer / issuer warps' waits

struct RBGeom {
    static constexpr int CA = 64, CH = 32, CB = 64;            // channels: in, hidden, out
    static constexpr int XBOX = 8 * 10 * 16;                    // one group box of x: [8 planes][10 px][16 B]
    static constexpr int XROW = 16 * XBOX;                      // 20480 B per x row
    static constexpr int XSTAGES = RB_XSTAGES;
    static constexpr int HPLANE = 10 * 16;
    static constexpr int HBOX = 4 * HPLANE;                     // hidden group box: [4 planes][10 px][16 B]
    static constexpr int HROW = 16 * HBOX;                      // 10240 B
    static constexpr int HSLOTS = RB_HSLOTS;
    static constexpr int KA = 3 * (CA / 16), KB = 3 * (CH / 16);           // 12, 6 K-steps per row
    static constexpr int WA_STEP = 2 * 3 * CH * 16, WB_STEP = 2 * 3 * CB * 16;   // 3072, 6144
    static constexpr int WA_BYTES = KA * WA_STEP, WB_BYTES = KB * WB_STEP;       // 36864 each
    static constexpr int RH = RB_RH, RY = RB_RY;                        // TMEM accumulator rings
    static constexpr int KSPLIT = RB_KSPLIT ? 2 : 1;                   // convA issuers = partial hidden accumulator rings
    static constexpr int TM_H = 0, TM_H1 = RH * CH, TM_Y = KSPLIT * RH * CH;   // column bases
    static constexpr int OFF_WA = 0, OFF_WB = WA_BYTES, OFF_X = OFF_WB + WB_BYTES, OFF_H = OFF_X + XSTAGES * XROW;
    static constexpr int OFF_BIAS = OFF_H + HSLOTS * HROW, OFF_BAR = OFF_BIAS + (CH + CB) * 4;
    static constexpr int SMEM_BYTES = 1024 + OFF_BAR + 512;
    // ---- CTA-pair variant (cta_group::2): every CTA holds HALF of each B piece.  A folded MMA covers slices [s0, s0 + n) of
    // B = [W(dy=2) | W(dy=1) | W(dy=0)]; the hardware takes the first N/2 rows of a piece from the leader CTA and the second
    // N/2 from its peer, both at the SAME shared-memory address, so every piece that is ever issued needs its own placement:
    //   arrangement 0: piece (0,3)                  the common case: a whole three-row window
    //   arrangement 1: pieces (0,1) then (1,2)      a window split after its first row by the ring's wrap-around / a top edge
    //   arrangement 2: pieces (0,2) then (2,1)      ... split after its second row / a bottom edge
    //   arrangement 3: piece (1,1)                  one-row segments
    // rows per CTA and arrangement: 3*CO/2 (arrangement 3: CO/2); bytes: K-steps x 2 granules x rows x 16.
    static constexpr int XSTAGES2 = RB_XSTAGES2;
    static constexpr int W2_ARR = KA * 2 * (3 * CH / 2) * 16;                        // 18432 (convB: KB * 2 * (3*CB/2) * 16 = the same)
    static constexpr int W2_P3 = KA * 2 * (CH / 2) * 16;                             // 6144
    static constexpr int W2_BYTES = 3 * W2_ARR + W2_P3;                              // 61440 per conv and CTA
    static_assert(KB * 2 * (3 * CB / 2) * 16 == W2_ARR && KB * 2 * (CB / 2) * 16 == W2_P3, "convA / convB pair packs have one size");
    static constexpr int OFF2_WA = 0, OFF2_WB = W2_BYTES, OFF2_X = 2 * W2_BYTES, OFF2_H = OFF2_X + XSTAGES2 * XROW;
    static constexpr int OFF2_BIAS = OFF2_H + HSLOTS * HROW, OFF2_BAR = OFF2_BIAS + (CH + CB) * 4;
    static constexpr int SMEM2_BYTES = 1024 + OFF2_BAR + 512;
    static_assert(SMEM2_BYTES <= 227 * 1024, "shared memory budget (pair)");
    static constexpr int THREADS = 640;            // 4 control warps (one spare) + 8 hidden-epilogue + 8 output-epilogue warps
    static_assert(KSPLIT * RH * CH + RY * CB == 512, "the accumulator rings fill TMEM");
    static_assert(KSPLIT == 1 || RB_BIAS_MODE == 1, "the K-split hidden rings are armed with zero");
    static_assert(RH >= 4 && RY >= 4, "a ring holds three accumulating rows and one draining");
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
    static_assert(XSTAGES <= 8 && RH <= 8 && HSLOTS <= 4 && RY <= 8, "barrier array layout");
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
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_zero16(uint32_t taddr) {
    const uint32_t z = 0u;
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};"
        ::"r"(taddr), "r"(z) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_zero32(uint32_t taddr) {   // one zero register, named 32 times
    const uint32_t z = 0u;
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};"
        ::"r"(taddr), "r"(z) : "memory");
}
__device__ __forceinline__ void sts_u4(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// L2 eviction priorities as cache-policy operands (the encodings createpolicy.fractional.L2::evict_{first,last} produce for a
// fraction of 1.0).  RB_L2_HINTS: a block's result is the next block's input — kept in L2 with evict_last when it is stored and
// released with evict_first on its last read (the residual add), so that the next launch finds most of its 134 MB input in the
// 126 MB L2 instead of reading rows that were written a whole launch (268 MB of traffic) earlier.
#ifndef RB_L2_HINTS
#define RB_L2_HINTS 3      // bit 0: evict_first on the residual load, bit 1: evict_last on the result store (forward 803.9 -> 800.7 us, profiles/ab_l2hints_r2.log)
#endif
constexpr uint64_t L2_EVICT_FIRST = 0x12F0000000000000ull, L2_EVICT_LAST = 0x14F0000000000000ull;
__device__ __forceinline__ uint4 rb_ld_res(const uint4* p) {
#if RB_L2_HINTS & 1
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p), "l"(L2_EVICT_FIRST));
    return r;
#elif RB_IO_NOALLOC >= 1
    return ldg_nc16(p);
#else
    return __ldg(p);
#endif
}
__device__ __forceinline__ void rb_st_out(uint4* p, uint4 v) {
#if RB_L2_HINTS & 2
    asm volatile("st.global.L1::no_allocate.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(L2_EVICT_LAST) : "memory");
#elif RB_IO_NOALLOC >= 2
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
#else
    *p = v;
#endif
}
__device__ __forceinline__ void rb_trace(const ResblockParams& p, int row, int k) {
    if (SRB_DBG(p.dbg_trace != nullptr) && blockIdx.x == 0 && row < 32) p.dbg_trace[row * 32 + k] = clock64();
}

// a segment: output rows [y0, y1) of strip-set ss, with the hidden / input row ranges it needs (clipped to the image)
struct RBSeg { int ss, y0, y1, hlo, hhi, xlo, xhi; };
// (ss_mul, ss_add: a CTA pair marches over PAIRS of strip-sets in lock-step — rank r of the pair owns strip-set 2 * q + r)
template <class Params>
__device__ __forceinline__ bool rb_next_seg(const Params& p, int& g, int g1, RBSeg& s, int ss_mul = 1, int ss_add = 0) {
    if (g >= g1) return false;
    const int q = g / p.H;
    s.ss = q * ss_mul + ss_add;
    s.y0 = g - q * p.H;
    int n = g1 - g;
    s.y1 = s.y0 + n < p.H ? s.y0 + n : p.H;
    s.hlo = s.y0 > 0 ? s.y0 - 1 : 0;
    s.hhi = s.y1 < p.H ? s.y1 : p.H - 1;
    s.xlo = s.hlo > 0 ? s.hlo - 1 : 0;
    s.xhi = s.hhi + 1 < p.H ? s.hhi + 1 : p.H - 1;
    g += s.y1 - s.y0;
    return true;
}

// Which 8-pixel group each of the 16 slots of strip-set ss is.
//   TS == 0 (GX <= 16): a tile holds TR = 16 / GX image rows, the same row of TR consecutive images; slots past TR * GX idle.
//   TS  > 0 (GX  > 16): a tile is 16 consecutive groups of one image row, an image row is TS strips advancing by 16 - 2 * OV
//   groups.  With OV = 1 (kernels whose result needs the neighbouring groups' data: the fused block's hidden halo, the
//   tail's dx exchange) the first / last slot of a strip only feeds its neighbour — its own result comes from the adjacent
//   strip — unless it is the image's edge group.
// `in`: the slot is a group of the image (else its operand rows are zero), `out`: this tile produces the group's result.
#ifndef RB_PF
#define RB_PF 0
#endif
// Narrow rows (TS == 0: TR image rows of GX groups side by side in a tile): slots in GROUP-major order (slot = gx * TR + image), so
// that ONE TMA box [10 px][8 planes][TR images] fills the TR slots of a group column: GX TMA instructions per x row instead of 16.
// Measured (profiles/rb_trace_tma_r2.log): the producer warp issues one TMA instruction per 70-90 clk, 16 per row put 1 400 clk between
// "stage free" and "row requested" and made the x ring (3 stages in the CTA-pair kernel) the loop that set the row period.
#ifndef RB_GM
#define RB_GM 1
#endif
#ifndef RB_EARLY_REL
#define RB_EARLY_REL 1
#endif
#ifndef RB_HALO_MERGED
#define RB_HALO_MERGED 1
#endif
#ifndef RB_IO_NOALLOC
#define RB_IO_NOALLOC 2      // 1: residual loads with L1::no_allocate; 2: the result's stores as well (forward 835.6 -> 833.5 us, profiles/ab_halo_io_r2.log)
#endif
struct RBSlot { int n, gx; bool in, out; };
template <class Params>
__device__ __forceinline__ RBSlot rb_slot(const Params& p, int ss, int j) {
    RBSlot s;
    if (p.TS == 0) {
        const int r = j / p.GX;
        s.n = ss * p.TR + r;
        s.gx = j - r * p.GX;
        s.in = r < p.TR && s.n < p.N;
        s.out = s.in;
    } else {
        s.n = ss / p.TS;
        s.gx = (ss - s.n * p.TS) * (16 - 2 * p.OV) + j;
        s.in = s.gx < p.GX && s.n < p.N;          // (n >= N: the phantom strip-set of a CTA pair when the count is odd)
        s.out = s.in && (p.OV == 0 || ((j > 0 || s.gx == 0) && (j < 15 || s.gx == p.GX - 1)));
    }
    if (!s.in) { s.n = p.N; s.gx = 0; }        // as TMA coordinates: outside the tensor, the box is zero-filled
    return s;
}

// fused residual block: narrow tiles in group-major slot order (RB_GM above); wide tiles as rb_slot
template <class Params>
__device__ __forceinline__ RBSlot rb_slot_gm(const Params& p, int ss, int j) {
    if (!RB_GM || p.TS != 0) return rb_slot(p, ss, j);
    RBSlot s;
    s.gx = j / p.TR;
    s.n = ss * p.TR + (j - s.gx * p.TR);
    s.in = s.gx < p.GX && s.n < p.N;
    s.out = s.in;
    if (!s.in) { s.n = p.N; s.gx = 0; }
    return s;
}

// Per-row issue plan of the folded MMAs: D slots of rows [c_lo, c_lo + nsl) (ring counters) += A * B[slice s0 ..],
// split in two at the ring's wrap-around.  Everything that does not depend on the K-step is computed once per row —
// the issuing thread is a single lane and its instruction latency is on the critical path.
struct RBPlan { uint32_t d0, d1, idesc0, idesc1, boff0, boff1; bool two; };
template <int CO, int R, bool BF16>
__device__ __forceinline__ RBPlan rb_plan(uint32_t tm_ring, int slot, int nsl, int s0) {
    RBPlan q;
    const int n1 = R - slot < nsl ? R - slot : nsl;
    q.d0 = tm_ring + slot * CO;
    q.idesc0 = umma_idesc_f16(128, n1 * CO, BF16 ? 1 : 0);
    q.boff0 = (uint32_t)((s0 * CO * 16) >> 4);
    q.two = n1 < nsl;
    q.d1 = tm_ring;
    q.idesc1 = umma_idesc_f16(128, (nsl - n1 > 0 ? nsl - n1 : 1) * CO, BF16 ? 1 : 0);
    q.boff1 = (uint32_t)(((s0 + n1) * CO * 16) >> 4);
    return q;
}
__device__ __forceinline__ void rb_issue(const RBPlan& q, uint64_t a_desc, uint64_t b_desc) {
    umma_f16_ss(q.d0, a_desc, b_desc + q.boff0, q.idesc0, 1u);
    if (q.two) umma_f16_ss(q.d1, a_desc, b_desc + q.boff1, q.idesc1, 1u);
}

// CTA-pair issue plan: the same window as RBPlan, every piece with its own B placement (see RBGeom) and M = 256
struct RBPlan2 { uint32_t d0, d1, idesc0, idesc1, ks0, ks1; uint64_t b0, b1; bool two; };
template <int CO, int KST, bool BF16>
__device__ __forceinline__ void rb_piece2(uint32_t w_addr, int s0, int n, uint64_t& b, uint32_t& ks, uint32_t& idesc) {
    constexpr int ROWS = 3 * CO / 2, ARR = KST * 2 * ROWS * 16;
    int arr, off;
    if (n == 3) { arr = 0; off = 0; }
    else if (n == 2) { arr = s0 == 1 ? 1 : 2; off = s0 == 1 ? CO / 2 : 0; }
    else { arr = s0 == 0 ? 1 : (s0 == 2 ? 2 : 3); off = s0 == 2 ? CO : 0; }
    const int rows = arr == 3 ? CO / 2 : ROWS;
    b = umma_desc_nosw(w_addr + arr * ARR + off * 16, rows * 16, 128);
    ks = (uint32_t)((2 * rows * 16) >> 4);
    idesc = umma_idesc_f16(256, n * CO, BF16 ? 1 : 0);
}
template <int CO, int R, int KST, bool BF16>
__device__ __forceinline__ RBPlan2 rb_plan2(uint32_t tm_ring, uint32_t w_addr, int slot, int nsl, int s0) {
    RBPlan2 q;
    const int n1 = R - slot < nsl ? R - slot : nsl;
    q.d0 = tm_ring + slot * CO;
    rb_piece2<CO, KST, BF16>(w_addr, s0, n1, q.b0, q.ks0, q.idesc0);
    q.two = n1 < nsl;
    q.d1 = tm_ring;
    q.b1 = q.b0; q.ks1 = q.ks0; q.idesc1 = q.idesc0;
    if (q.two) rb_piece2<CO, KST, BF16>(w_addr, s0 + n1, nsl - n1, q.b1, q.ks1, q.idesc1);
    return q;
}
__device__ __forceinline__ void rb_issue2(const RBPlan2& q, uint64_t a_desc, int kk) {
    umma_f16_ss_pair(q.d0, a_desc, q.b0 + (uint64_t)(kk * q.ks0), q.idesc0, 1u);
    if (q.two) umma_f16_ss_pair(q.d1, a_desc, q.b1 + (uint64_t)(kk * q.ks1), q.idesc1, 1u);
}

// ---- Steady-state issue path of the CTA-pair kernel -------------------------------------------------------------------------
// ncu's samples (profiles/ncu_fwd_r2_ingest_full.txt, source page) have the MMA-issuing threads busy three quarters of the time
// with their own instruction stream — about 270 SASS instructions per row and issuer, most of them the per-row plan (which
// arrangement, which piece, 64-bit descriptor arithmetic) — while the UTCHMMAs themselves almost never stall.  For a window of
// three rows with no clipping (every row but a segment's first and last) the plan depends only on the ring slot of the window's
// first row, so it is compiled four times with everything folded: a descriptor is {low word = a run-time base + an immediate,
// high word = a constant}, one add per operand and MMA.  Same MMAs in the same order as rb_plan2 / rb_issue2.
#ifndef RB_STEADY
#define RB_STEADY 1
#endif
template <int CO, int KST, int S0, int N, bool BF16>
struct RBPieceC {          // rb_piece2 at compile time
    static constexpr int ROWS = 3 * CO / 2, ARR = KST * 2 * ROWS * 16;
    static constexpr int arr = N == 3 ? 0 : N == 2 ? (S0 == 1 ? 1 : 2) : (S0 == 0 ? 1 : (S0 == 2 ? 2 : 3));
    static constexpr int off = N == 3 ? 0 : N == 2 ? (S0 == 1 ? CO / 2 : 0) : (S0 == 2 ? CO : 0);
    static constexpr int rows = arr == 3 ? CO / 2 : ROWS;
    static constexpr uint32_t lo = (uint32_t)((arr * ARR + off * 16) >> 4) + ((uint32_t)rows << 16);    // + (w_addr >> 4): address and LBO fields
    static constexpr uint32_t ks = (uint32_t)((2 * rows * 16) >> 4);
    static constexpr uint32_t idesc = umma_idesc_f16(256, N * CO, BF16 ? 1 : 0);
};
__device__ __forceinline__ void umma_pair_lohi(uint32_t d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc) {
    asm volatile(
        "{\n\t.reg .b64 da, db;\n\t.reg .pred p;\n\t"
        "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\tsetp.eq.u32 p, 0, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n\t}"
        ::"r"(d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc) : "memory");
}
// NK K-steps i: the A operand of step i starts (i & 1) * 20 + (i >> 1) granules into the row (channel pair, dx), its filters are
// K-step (CONVA ? (i >> 1) * 4 + (i & 1) : i) of the pack; a K-split issuer's channel offset is part of alo / wlo
template <int CO, int R, int KST, int SLOT, int NK, bool CONVA, bool BF16>
__device__ __forceinline__ void rb_steady2(uint32_t tm_ring, uint32_t wlo, uint32_t alo, uint32_t ahi) {
    constexpr int N1 = (R - SLOT < 3) ? R - SLOT : 3;
    constexpr bool TWO = N1 < 3;
    using Q0 = RBPieceC<CO, KST, 0, N1, BF16>;
    using Q1 = RBPieceC<CO, KST, TWO ? N1 : 0, TWO ? 3 - N1 : 3, BF16>;
    constexpr uint32_t BHI = (128u >> 4) | (1u << 14);
#pragma unroll
    for (int i = 0; i < NK; ++i) {
        const uint32_t aimm = (uint32_t)((i & 1) * 20 + (i >> 1));
        const uint32_t kimm = (uint32_t)(CONVA ? (i >> 1) * 4 + (i & 1) : i);
        umma_pair_lohi(tm_ring + SLOT * CO, alo + aimm, ahi, wlo + Q0::lo + kimm * Q0::ks, BHI, Q0::idesc);
        if constexpr (TWO) umma_pair_lohi(tm_ring, alo + aimm, ahi, wlo + Q1::lo + kimm * Q1::ks, BHI, Q1::idesc);
    }
}

template <bool BF16, bool PAIR = false>
__global__ void __launch_bounds__(RBGeom::THREADS, 1) resblock_kernel(const __grid_constant__ CUtensorMap tmap_x,
                                                                      const ResblockParams p) {
    using G = RBGeom;
    extern __shared__ uint8_t smem_raw[];
    // (offset arithmetic on the __shared__ array, not an integer round trip: the compiler keeps the shared address space
    // and the epilogues' plain loads / stores become LDS / STS instead of generic accesses)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    constexpr int XST = PAIR ? G::XSTAGES2 : G::XSTAGES;
    uint8_t* s_wA = smem + (PAIR ? G::OFF2_WA : G::OFF_WA);
    uint8_t* s_wB = smem + (PAIR ? G::OFF2_WB : G::OFF_WB);
    uint8_t* s_x = smem + (PAIR ? G::OFF2_X : G::OFF_X);
    uint8_t* s_h = smem + (PAIR ? G::OFF2_H : G::OFF_H);
    float* s_bias = reinterpret_cast<float*>(smem + (PAIR ? G::OFF2_BIAS : G::OFF_BIAS));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (PAIR ? G::OFF2_BAR : G::OFF_BAR));
    // CTA pair: rank 0 (the leader) issues every MMA for both CTAs; the consumer -> issuer barriers (haempty, hsfull, yaempty,
    // and xfull for the TMA bytes) live in the LEADER and collect the arrivals of both CTAs, the issuer -> consumer barriers
    // exist in both CTAs and are signalled by multicast commits
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const bool leader = rank == 0u;
    const uint32_t bars_leader = PAIR ? mapa_u32(smem_u32(bars), 0u) : 0u;      // shared::cluster address of the leader's barrier array
    uint64_t* xfull = bars;            // [XSTAGES]  TMA -> MMA
    uint64_t* xempty = bars + 8;       // [XSTAGES]  MMA -> TMA
    uint64_t* hafull = bars + 16;      // [RH] hidden accumulator row complete        MMA -> hidden epilogue
    uint64_t* haempty = bars + 24;     // [RH] drained and re-zeroed                  hidden epilogue -> MMA
    uint64_t* hsfull = bars + 32;      // [HSLOTS] hidden row written to shared memory   hidden epilogue -> MMA
    uint64_t* hsempty = bars + 36;     // [HSLOTS] consumed by convB                     MMA -> hidden epilogue
    uint64_t* yafull = bars + 40;      // [RY] output accumulator row complete        MMA -> output epilogue
    uint64_t* yaempty = bars + 48;     // [RY] drained and re-zeroed                  output epilogue -> MMA
    uint64_t* wbar = bars + 56;
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 57);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (SRB_DBG(p.dbg_trace != nullptr) && threadIdx.x == 0) {           // per-CTA timing: [1024 + 8 * cta + k]
        long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        p.dbg_trace[1024 + 8 * blockIdx.x] = gt;
        p.dbg_trace[1024 + 8 * blockIdx.x + 1] = clock64();
        // per-launch records [2304 + 4 * launch]: earliest CTA start, latest CTA start, earliest end, latest end (ns)
        const long long old = (long long)atomicAdd((unsigned long long*)&p.dbg_trace[2300], 1ull);
        const int launch = (int)(old / gridDim.x);
        if (launch < 400) {
            atomicMin(&p.dbg_trace[2304 + 4 * launch], gt);
            atomicMax(&p.dbg_trace[2304 + 4 * launch + 1], gt);
        }
        p.dbg_trace[1024 + 8 * blockIdx.x + 6] = launch;
    }

    if SRB_DBG(p.dbg_skip & 4096) {        // launch-overhead probe: nothing but the stamps
        if (SRB_DBG(p.dbg_trace != nullptr) && threadIdx.x == 0) {
            long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
            p.dbg_trace[1024 + 8 * blockIdx.x + 5] = gt;
            const int launch = (int)p.dbg_trace[1024 + 8 * blockIdx.x + 6];
            if (launch < 400) {
                atomicMin(&p.dbg_trace[2304 + 4 * launch + 2], gt);
                atomicMax(&p.dbg_trace[2304 + 4 * launch + 3], gt);
            }
        }
        return;
    }
    // zero the hidden ring once: halo slots at image borders are never written and must read as zero padding
    for (int i = threadIdx.x; i < G::HSLOTS * G::HROW / 16; i += G::THREADS)
        reinterpret_cast<uint4*>(s_h)[i] = make_uint4(0, 0, 0, 0);
    if (RB_GM && p.TS == 0 && p.GX * p.TR < 16)      // slots no TMA box ever fills: their operand rows read as zeros
        for (int i = threadIdx.x; i < XST * G::XROW / 16; i += G::THREADS) reinterpret_cast<uint4*>(s_x)[i] = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < G::CH + G::CB; i += G::THREADS) s_bias[i] = p.bias[i];
    if (threadIdx.x == 0) {
        constexpr int NC = PAIR ? 2 : 1;     // CTAs whose warps arrive on the leader's consumer -> issuer barriers
        for (int i = 0; i < XST; ++i) { mbar_init(&xfull[i], NC); mbar_init(&xempty[i], G::KSPLIT); }
        for (int i = 0; i < G::RH; ++i) { mbar_init(&hafull[i], G::KSPLIT); mbar_init(&haempty[i], 4 * NC); }
        for (int i = 0; i < G::HSLOTS; ++i) { mbar_init(&hsfull[i], 4 * NC); mbar_init(&hsempty[i], 1); }
        for (int i = 0; i < G::RY; ++i) { mbar_init(&yafull[i], 1); mbar_init(&yaempty[i], 8 * NC); }
        mbar_init(wbar, 1);
        fence_mbar_init();
        tma_prefetch_desc(&tmap_x);
        if constexpr (PAIR) {                // this CTA's half of the filters (the leader's issuers need both halves in place
            fence_proxy_async_smem();        // before their first MMA: waited for below, ahead of the cluster barrier)
            mbar_arrive_expect_tx(wbar, 2 * G::W2_BYTES);
            const uint8_t* ga = p.wA2 + (size_t)rank * G::W2_BYTES;
            const uint8_t* gb = p.wB2 + (size_t)rank * G::W2_BYTES;
            for (int off = 0; off < G::W2_BYTES; off += 12288) bulk_load_1d(s_wA + off, ga + off, 12288, wbar);
            for (int off = 0; off < G::W2_BYTES; off += 12288) bulk_load_1d(s_wB + off, gb + off, 12288, wbar);
        }
    }
    fence_proxy_async_smem();
    if (warp == 0) { __syncwarp(); if constexpr (PAIR) tmem_alloc_pair(s_tmem, 512); else tmem_alloc(s_tmem, 512); }
    tc_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    if (warp >= 4 && warp < 8) {
        // every MMA of this kernel accumulates, and an accumulator row starts at its bias (so the epilogues do not add
        // it): the rings are armed here once and re-armed by the epilogue that drains a row
        const uint32_t lg = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll 1
        for (int c = 0; c < 512; c += 32) {
            const int b0 = c < G::TM_Y ? 0 : G::CH + ((c - G::TM_Y) & 32);
            uint32_t bv[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) bv[k] = RB_BIAS_MODE == 0 ? __float_as_uint(p.bias[b0 + k]) : 0u;
            tmem_st32(tmem_base + lg + c, bv);
        }
        tmem_st_wait();
    }
    if constexpr (PAIR) { if (warp == 0) mbar_wait(wbar, 0, 32); }
    tc_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
    tc_fence_after();

    if (SRB_DBG(p.dbg_trace != nullptr) && threadIdx.x == 0) p.dbg_trace[1024 + 8 * blockIdx.x + 2] = clock64();
    // Everything above touched only this launch's constants.  The next launch in the stream may now start its own
    // prologue on every SM this grid vacates; the warps below that touch activations first wait for the previous launch.
    pdl_launch_dependents();
    // (pair: total_rows counts rows of strip-set PAIRS; both CTAs of a cluster take the same run and march in lock-step)
    const int n_units = PAIR ? (int)gridDim.x >> 1 : (int)gridDim.x, unit = PAIR ? (int)blockIdx.x >> 1 : (int)blockIdx.x;
    const int g_begin = (int)((long long)unit * p.total_rows / n_units);
    const int g_end = SRB_DBG(p.dbg_skip & 2048) ? g_begin : (int)((long long)(unit + 1) * p.total_rows / n_units);   // 2048: prologue-only probe
    const int ss_mul = PAIR ? 2 : 1, ss_add = (int)rank;
    const size_t HW = (size_t)p.H * p.W;

    if (warp == 0) {
        // ===================== TMA producer: filters once, then one x row (16 group boxes) per stage =====================
        if (!PAIR && elect_one()) {
            mbar_arrive_expect_tx(wbar, G::WA_BYTES + G::WB_BYTES);
            for (int off = 0; off < G::WA_BYTES; off += 12288) bulk_load_1d(s_wA + off, p.wA + off, 12288, wbar);
            for (int off = 0; off < G::WB_BYTES; off += 12288) bulk_load_1d(s_wB + off, p.wB + off, 12288, wbar);
        }
        __syncwarp();
        pdl_wait();                          // x is the previous launch's output
        const bool gm = RB_GM && p.TS == 0;
        const int gm_box = p.TR * G::XBOX;
        const uint32_t xbytes = gm ? (uint32_t)(p.GX * gm_box) : (uint32_t)G::XROW;
        int xc = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg, ss_mul, ss_add)) {
            const RBSlot sl = rb_slot_gm(p, sg.ss, lane & 15);
            const int n = sl.n, gx = sl.gx;
            // group-major tiles: lane gx loads the box of group column gx for the tile's TR images (images beyond N: zero-filled)
            const int gm_n0 = sg.ss * p.TR < p.N ? sg.ss * p.TR : p.N;
#pragma unroll 1
            for (int r = sg.xlo; r <= sg.xhi; ++r, ++xc) {
                const int st = xc % XST;
                if constexpr (RB_PF > 0) {
                    // L2 prefetch of the x row RB_PF steps ahead (one 128-byte line per slot and plane): the TMA boxes of that row
                    // then hit L2 instead of waiting on HBM with only XST stages of requests in flight
                    const int rr = r + RB_PF;
                    if (lane < 16 && sl.in && rr < p.H) {
                        const char* a = (const char*)p.x + (((size_t)n * 8 * p.H + rr) * p.W + (size_t)gx * 8) * 16;
#pragma unroll
                        for (int pl = 0; pl < 8; ++pl) prefetch_l2(a + (size_t)pl * HW * 16);
                    }
                }
                mbar_wait<RB_CTRL>(&xempty[st], (uint32_t)(((xc / XST) & 1) ^ 1), 31);
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0) rb_trace(p, xc, 19);         // x stage free
                if constexpr (PAIR) {
                    // both CTAs' rows of this step complete the LEADER's barrier (two arrivals, 2 x XROW bytes)
                    const uint32_t xf = bars_leader + (uint32_t)(st * 8);
                    if (lane == 0) mbar_arrive_expect_tx_cluster(xf, xbytes);
                    __syncwarp();
                    if (gm) {
                        if (lane < p.GX)
                            tma_load_4d_pair(s_x + st * G::XROW + lane * gm_box, &tmap_x, xf, (lane * 8 - 1) * 4, r, 0, gm_n0);
                    } else if (lane < 16)
                        tma_load_4d_pair(s_x + st * G::XROW + lane * G::XBOX, &tmap_x, xf, (gx * 8 - 1) * 4, r, 0, n);
                    __syncwarp();
                    if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0) rb_trace(p, xc, 20);         // x row requested
                    continue;
                }
                if SRB_DBG(p.dbg_skip & 1) { if (lane == 0) mbar_arrive(&xfull[st]); __syncwarp(); continue; }
                if (lane == 0) mbar_arrive_expect_tx(&xfull[st], xbytes);
                __syncwarp();
                if (gm) {
                    if (lane < p.GX)
                        tma_load_4d(s_x + st * G::XROW + lane * gm_box, &tmap_x, &xfull[st], (lane * 8 - 1) * 4, r, 0, gm_n0);
                } else if (lane < 16)
                    tma_load_4d(s_x + st * G::XROW + lane * G::XBOX, &tmap_x, &xfull[st], (gx * 8 - 1) * 4, r, 0, n);
                __syncwarp();
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0) rb_trace(p, xc, 20);         // x row requested
            }
        }
    } else if ((warp == 1 || (G::KSPLIT == 2 && warp == 3)) && leader) {
        // ===================== convA MMA issuer(s): one x row per step, N = 3 x 32 into the hidden accumulator ring =====================
        // K split: warp 1 owns input channels 0..31 (K-steps c2 = 0, 1 of every dx) and ring H0, warp 3 channels 32..63 and H1
        const int kh = warp == 3 ? 1 : 0;
        const uint32_t tm_ring_a = tmem_base + (kh ? G::TM_H1 : G::TM_H);
        // (convA and convB are issued by two different warps: a single thread cannot prepare 18 MMAs + 5 commits per row
        // fast enough, and the tensor pipe interleaves the two warps' independent accumulation chains by itself)
        const uint64_t adescA0 = umma_desc_nosw(0, 160, G::XBOX);           // x row: planes 160 B apart, groups XBOX apart
        const uint64_t bdescA0 = umma_desc_nosw(smem_u32(s_wA), 3 * G::CH * 16, 128);
        if constexpr (!PAIR) mbar_wait(wbar, 0, 32);
        const uint32_t wA_u32 = smem_u32(s_wA);
        // ring positions are tracked incrementally (no divisions on the issue path)
        int xst = 0; uint32_t xph = 0;                 // x stage / phase of the next x row
        int hslot0 = 0;                                // ring slot of the segment's first hidden row (hidden counter hbase)
        int acq_slot = 0; uint32_t acq_ph = 1;         // next hidden slot to acquire, parity to wait for (empty: phase ^ 1)
        int hbase = 0, h_acq = 0, g = g_begin, xc = 0;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg, ss_mul, ss_add)) {
#pragma unroll 1
            for (int r = sg.xlo; r <= sg.xhi; ++r, ++xc) {
                mbar_wait<RB_CTRL>(&xfull[xst], xph, 33);
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0 && kh == 0) rb_trace(p, xc, 21);         // x row arrived
                const int lo = r - 1 > sg.hlo ? r - 1 : sg.hlo, hi = r + 1 < sg.hhi ? r + 1 : sg.hhi;
                const int a_lo = hbase + lo - sg.hlo, a_n = hi - lo + 1, a_s0 = lo - (r - 1);
                for (; h_acq < a_lo + a_n; ++h_acq) {
                    mbar_wait<RB_CTRL>(&haempty[acq_slot], acq_ph, 34);
                    if (++acq_slot == G::RH) { acq_slot = 0; acq_ph ^= 1; }
                }
                const int slot_lo = (hslot0 + (lo - sg.hlo)) % G::RH;
                tc_fence_after();
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0 && kh == 0) rb_trace(p, xc, 0);          // waits of this x row satisfied
                if (elect_one()) {
                    const uint64_t adA = adescA0 + (uint64_t)(smem_u32(s_x + xst * G::XROW) >> 4);
                    if constexpr (PAIR) {
                        if (RB_STEADY && G::KSPLIT == 2 && G::RH == 4 && a_n == 3 && a_s0 == 0) {
                            // (K-split: this issuer's channel pairs c2 = 2 * kh + cc start kh * 40 granules into the row, its K-steps at 2 * kh)
                            const uint32_t alo = (10u << 16) + (smem_u32(s_x + xst * G::XROW) >> 4) + (uint32_t)kh * 40u;
                            const uint32_t wlo = (wA_u32 >> 4) + (uint32_t)kh * 2u * RBPieceC<G::CH, G::KA, 0, 3, BF16>::ks;
                            constexpr uint32_t AHI = (uint32_t)(G::XBOX >> 4) | (1u << 14);
                            // (a full window: hidden row r - 1 sits in ring slot slot_lo and is now complete; row r is not the segment's last)
#define RB_CASE_A(S) case S: rb_steady2<G::CH, 4, G::KA, S, 6, true, BF16>(tm_ring_a, wlo, alo, AHI); umma_commit_pair(&xempty[xst]); umma_commit_pair(&hafull[S]); break;
                            switch (slot_lo) { RB_CASE_A(0) RB_CASE_A(1) RB_CASE_A(2) default: RB_CASE_A(3) }
#undef RB_CASE_A
                        } else {
                        const RBPlan2 q2 = rb_plan2<G::CH, G::RH, G::KA, BF16>(tm_ring_a, wA_u32, slot_lo, a_n, a_s0);
#pragma unroll
                        for (int dx = 0; dx < 3; ++dx)
#pragma unroll
                            for (int cc = 0; cc < (G::KSPLIT == 2 ? 2 : 4); ++cc) {
                                const int c2 = (G::KSPLIT == 2 ? 2 * kh : 0) + cc, kk = dx * 4 + c2;
                                rb_issue2(q2, adA + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4), kk);
                            }
                        umma_commit_pair(&xempty[xst]);
                        if (r - 1 >= sg.hlo) umma_commit_pair(&hafull[(hslot0 + (r - 1 - sg.hlo)) % G::RH]);
                        if (r == sg.xhi && r <= sg.hhi) umma_commit_pair(&hafull[(hslot0 + (r - sg.hlo)) % G::RH]);
                        }
                    } else {
                    const RBPlan q = rb_plan<G::CH, G::RH, BF16>(tm_ring_a, SRB_DBG(p.dbg_skip & 512) && slot_lo > G::RH - 3 ? G::RH - 3 : slot_lo, a_n, a_s0);
                    if (!SRB_DBG(p.dbg_skip & (2 | 16))) {
                        if constexpr (G::KSPLIT == 2) {
#pragma unroll
                            for (int dx = 0; dx < 3; ++dx)
#pragma unroll
                                for (int cc = 0; cc < 2; ++cc) {
                                    const int c2 = 2 * kh + cc, kk = dx * 4 + c2;
                                    rb_issue(q, adA + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4), bdescA0 + (uint64_t)((kk * G::WA_STEP) >> 4));
                                }
                        } else {
#pragma unroll
                            for (int kk = 0; kk < G::KA; ++kk) {
                                const int dx = kk >> 2, c2 = kk & 3;
                                rb_issue(q, adA + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4), bdescA0 + (uint64_t)((kk * G::WA_STEP) >> 4));
                            }
                        }
                    }
                    umma_commit(&xempty[xst]);
                    // hidden row r-1 is complete (all three input rows seen); at the last x row also row r itself
                    if (r - 1 >= sg.hlo) umma_commit(&hafull[(hslot0 + (r - 1 - sg.hlo)) % G::RH]);
                    if (r == sg.xhi && r <= sg.hhi) umma_commit(&hafull[(hslot0 + (r - sg.hlo)) % G::RH]);
                    }
                }
                __syncwarp();
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0 && kh == 0) rb_trace(p, xc, 1);          // MMAs + commits issued
                if (++xst == XST) { xst = 0; xph ^= 1; }
            }
            const int nh = sg.hhi - sg.hlo + 1;
            hbase += nh;
            hslot0 = (hslot0 + nh) % G::RH;
        }
    } else if (warp == 2 && leader) {
        // ===================== convB MMA issuer: one hidden row per step, N = 3 x 64 into the output accumulator ring =====================
        const uint64_t adescB0 = umma_desc_nosw(0, 160, G::HBOX);
        const uint64_t bdescB0 = umma_desc_nosw(smem_u32(s_wB), 3 * G::CB * 16, 128);
        if constexpr (!PAIR) mbar_wait(wbar, 0, 37);
        const uint32_t wB_u32 = smem_u32(s_wB);
        int hs = 0; uint32_t hph = 0;                  // hidden smem slot / phase of the next hidden row
        int yslot0 = 0;                                // ring slot of the segment's first output row
        int acq_slot = 0; uint32_t acq_ph = 1;
        int ybase = 0, ybase_h = 0, y_acq = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg, ss_mul, ss_add)) {
#pragma unroll 1
            for (int h = sg.hlo; h <= sg.hhi; ++h) {
                const int hc = ybase_h + h - sg.hlo;               // running hidden-row counter (trace only)
                mbar_wait<RB_CTRL>(&hsfull[hs], hph, 35);
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0) rb_trace(p, hc, 16);         // hidden row in shared memory
                const int lo = h - 1 > sg.y0 ? h - 1 : sg.y0, hi = h + 1 < sg.y1 - 1 ? h + 1 : sg.y1 - 1;
                const int b_lo = ybase + lo - sg.y0, b_n = hi - lo + 1, b_s0 = lo - (h - 1);
                for (; y_acq < b_lo + b_n; ++y_acq) {
                    mbar_wait<RB_CTRL>(&yaempty[acq_slot], acq_ph, 36);
                    if (++acq_slot == G::RY) { acq_slot = 0; acq_ph ^= 1; }
                }
                const int slot_lo = (yslot0 + (lo - sg.y0)) % G::RY;
                tc_fence_after();
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0) rb_trace(p, hc, 17);         // output accumulator rows acquired
                if (elect_one()) {
                    const uint64_t adB = adescB0 + (uint64_t)(smem_u32(s_h + hs * G::HROW) >> 4);
                    if constexpr (PAIR) {
                        if (RB_STEADY && G::RY == 4 && b_n == 3 && b_s0 == 0) {
                            const uint32_t alo = (10u << 16) + (smem_u32(s_h + hs * G::HROW) >> 4);
                            const uint32_t wlo = wB_u32 >> 4;
                            constexpr uint32_t AHI = (uint32_t)(G::HBOX >> 4) | (1u << 14);
                            const uint32_t tm_y = tmem_base + G::TM_Y;
                            // (a full window: output row h - 1 sits in ring slot slot_lo and is now complete; row h is not the segment's last)
#define RB_CASE_B(S) case S: rb_steady2<G::CB, 4, G::KB, S, G::KB, false, BF16>(tm_y, wlo, alo, AHI); umma_commit_pair(&hsempty[hs]); umma_commit_pair(&yafull[S]); break;
                            switch (slot_lo) { RB_CASE_B(0) RB_CASE_B(1) RB_CASE_B(2) default: RB_CASE_B(3) }
#undef RB_CASE_B
                        } else {
                        const RBPlan2 q2 = rb_plan2<G::CB, G::RY, G::KB, BF16>(tmem_base + G::TM_Y, wB_u32, slot_lo, b_n, b_s0);
#pragma unroll
                        for (int k = 0; k < G::KB; ++k) {
                            const int dx = k >> 1, c2 = k & 1;
                            rb_issue2(q2, adB + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4), k);
                        }
                        umma_commit_pair(&hsempty[hs]);
                        if (h - 1 >= sg.y0) umma_commit_pair(&yafull[(yslot0 + (h - 1 - sg.y0)) % G::RY]);
                        if (h == sg.hhi && h <= sg.y1 - 1) umma_commit_pair(&yafull[(yslot0 + (h - sg.y0)) % G::RY]);
                        }
                    } else {
                    const RBPlan q = rb_plan<G::CB, G::RY, BF16>(tmem_base + G::TM_Y, SRB_DBG(p.dbg_skip & 512) && slot_lo > G::RY - 3 ? G::RY - 3 : slot_lo, b_n, b_s0);
                    if (!SRB_DBG(p.dbg_skip & (2 | 32))) {
#pragma unroll
                        for (int k = 0; k < G::KB; ++k) {
                            const int dx = k >> 1, c2 = k & 1;
                            rb_issue(q, adB + (uint64_t)((2 * c2 * 160 + dx * 16) >> 4), bdescB0 + (uint64_t)((k * G::WB_STEP) >> 4));
                        }
                    }
                    umma_commit(&hsempty[hs]);
                    if (h - 1 >= sg.y0) umma_commit(&yafull[(yslot0 + (h - 1 - sg.y0)) % G::RY]);
                    if (h == sg.hhi && h <= sg.y1 - 1) umma_commit(&yafull[(yslot0 + (h - sg.y0)) % G::RY]);
                    }
                }
                __syncwarp();
                if (SRB_DBG(p.dbg_trace != nullptr) && lane == 0) rb_trace(p, hc, 18);         // convB MMAs + commits issued
                if (++hs == G::HSLOTS) { hs = 0; hph ^= 1; }
            }
            ybase_h += sg.hhi - sg.hlo + 1;
            const int ny = sg.y1 - sg.y0;
            ybase += ny;
            yslot0 = (yslot0 + ny) % G::RY;
        }
    }
    else if (warp >= 4 && warp < 12) {
        // ===================== hidden epilogue: TMEM -> bias, LeakyReLU -> 16-bit -> shared memory (convB's A operand) =====================
        // Two sets of 4 warps alternate rows: a row's hand-off time is dominated by fixed latencies (TMEM load, shared
        // stores, fences, barrier round trips), so rows are overlapped rather than split.
        const int lg = warp & 3;
        const int set = (warp - 4) >> 2;
        const int L = lg * 32 + lane;              // accumulator lane = pixel of the row
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + G::TM_H + ((uint32_t)(lg * 32) << 16);
        const uint32_t s_h_u32 = smem_u32(s_h);
        const bool tr = SRB_DBG(p.dbg_trace != nullptr) && lane == 0 && lg == 0;
        int hbase = 0, g = g_begin;
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg, ss_mul, ss_add)) {
            const RBSlot sl = rb_slot_gm(p, sg.ss, grp);
            const bool valid = sl.in && sl.gx * 8 + i < p.W;
            // halo pixels go to the neighbouring slot when that is the neighbouring group of the same image row (an idle or
            // out-of-image slot of a wide tile writes the zeros its neighbour needs as padding)
            // (group-major narrow tiles: the same image row's neighbouring groups are TR slots away; idle slots have no neighbours)
            const bool gmt = RB_GM && p.TS == 0;
            const int nb = gmt ? p.TR * G::HBOX : G::HBOX;
            const bool left = i == 0 && (gmt ? sl.in && sl.gx > 0 : grp > 0 && (p.TS > 0 || sl.gx > 0));
            const bool right = i == 7 && (gmt ? sl.in && sl.gx < p.GX - 1 : grp < 15 && (p.TS > 0 || sl.gx < p.GX - 1));
            const int halo_off = left ? 9 * 16 - nb : nb;
#pragma unroll 1
            for (int h = sg.hlo; h <= sg.hhi; ++h) {
                const int c = hbase + h - sg.hlo;
                if ((c & 1) != set) continue;
                const int slot = c % G::RH, hs = c % G::HSLOTS;
                if (tr) rb_trace(p, c, 2);        // hidden epilogue starts waiting
                mbar_wait<RB_SLEEP>(&hafull[slot], (uint32_t)((c / G::RH) & 1), 41);
                if (tr) rb_trace(p, c, 3);        // hidden accumulator complete
                tc_fence_after();
                uint32_t rr[32];
                tmem_ld32(t_lane + slot * G::CH, rr);
                if constexpr (G::KSPLIT == 2) {             // second partial sum (input channels 32..63)
                    uint32_t r1[32];
                    tmem_ld32(t_lane + G::TM_H1 + slot * G::CH, r1);
                    tmem_ld_wait();
#pragma unroll
                    for (int k = 0; k < 32; ++k) rr[k] = __float_as_uint(__uint_as_float(rr[k]) + __uint_as_float(r1[k]));
                    if (!SRB_DBG(p.dbg_skip & 64)) tmem_zero32(t_lane + G::TM_H1 + slot * G::CH);
                } else tmem_ld_wait();
                if (RB_BIAS_MODE == 1) {
                    if (!SRB_DBG(p.dbg_skip & 64)) tmem_zero32(t_lane + slot * G::CH);
                } else if (!SRB_DBG(p.dbg_skip & 64)) {    // re-arm the accumulator row with the bias
                    uint32_t bv[32];
#pragma unroll
                    for (int k = 0; k < 32; ++k) bv[k] = __float_as_uint(s_bias[k]);
                    tmem_st32(t_lane + slot * G::CH, bv);
                }
                if constexpr (RB_EARLY_REL) {      // hand the accumulator row back as soon as it is drained and re-armed (not after the
                    tmem_st_wait();                // shared-memory stores below): the convA issuer two rows ahead is waiting for it
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if constexpr (PAIR) mbar_arrive_cluster(bars_leader + (uint32_t)((24 + slot) * 8));
                        else mbar_arrive(&haempty[slot]);
                    }
                }
                if (tr) rb_trace(p, c, 11);
                uint4 o[4];
#pragma unroll
                for (int pl = 0; pl < 4; ++pl) {
                    float v[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float tt = __uint_as_float(rr[8 * pl + k]) + (RB_BIAS_MODE == 1 ? p.bias[8 * pl + k] : 0.f);
                        v[k] = fmaxf(tt, tt * p.slope);
                    }
                    o[pl].x = pack2<BF16>(v[0], v[1]); o[pl].y = pack2<BF16>(v[2], v[3]);
                    o[pl].z = pack2<BF16>(v[4], v[5]); o[pl].w = pack2<BF16>(v[6], v[7]);
                    if (!valid) o[pl] = make_uint4(0, 0, 0, 0);   // outside the image the hidden activation is convB's zero padding
                }
                mbar_wait<RB_SLEEP>(&hsempty[hs], (uint32_t)(((c / G::HSLOTS) & 1) ^ 1), 42);
                if (tr) rb_trace(p, c, 12);
                const uint32_t box = s_h_u32 + hs * G::HROW + grp * G::HBOX;
#pragma unroll
                for (int pl = 0; pl < 4; ++pl) {
                    sts_u4(box + pl * 160 + (i + 1) * 16, o[pl]);
#if RB_HALO_MERGED
                    // the edge pixels' copies into the neighbouring groups' halo slots — pixel 0 into the right halo slot of the group
                    // to the left, pixel 7 into the left halo slot of the group to the right — as ONE store: a 16-byte store is issued
                    // per quarter-warp, so two predicated stores with one active lane per quarter each cost 4 + 4 wavefronts of the L1
                    // data pipe (ncu source page: 3.5 per instruction), the merged one 4 (the two lanes of a quarter hit different banks)
                    if (left || right) sts_u4(box + halo_off + pl * 160, o[pl]);
#else
                    if (left) sts_u4(box - nb + pl * 160 + 9 * 16, o[pl]);           // right halo slot of the group to the left
                    if (right) sts_u4(box + nb + pl * 160, o[pl]);                   // left halo slot of the group to the right
#endif
                }
                if (tr) rb_trace(p, c, 13);
                if constexpr (!RB_EARLY_REL) { tmem_st_wait(); tc_fence_before(); }
                if (tr) rb_trace(p, c, 14);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) {
                    if constexpr (PAIR) {
                        if constexpr (!RB_EARLY_REL) mbar_arrive_cluster(bars_leader + (uint32_t)((24 + slot) * 8));
                        mbar_arrive_cluster(bars_leader + (uint32_t)((32 + hs) * 8));
                    } else {
                        if constexpr (!RB_EARLY_REL) mbar_arrive(&haempty[slot]);
                        mbar_arrive(&hsfull[hs]);
                    }
                }
                if (tr) rb_trace(p, c, 4);        // hidden row published
            }
            hbase += sg.hhi - sg.hlo + 1;
        }
    } else if (warp >= 12) {
        // ===================== output epilogue: 8 warps per row (two channel halves); + x (residual) -> y =====================
        // Residual in and result out go straight between global memory and registers: a warp's 32 lanes are 32
        // consecutive pixels, so each of a thread's four 16-byte plane granules is part of a 512-byte contiguous run.
        // (Staging the row in shared memory for TMA costs 64 KB of shared-memory traffic per row — residual in, read,
        // result written, store read-out — and this kernel is bound by the 128 B/clk shared-memory port: the MMAs alone
        // read 147 KB of operands per row.)  The residual of the next row is requested before this row's math.
        const int lg = warp & 3;
        const int half = (warp - 12) >> 2;         // channels 32*half .. : planes 4*half .. 4*half+3
        const int L = lg * 32 + lane;
        const int grp = L >> 3, i = L & 7;
        const uint32_t t_lane = tmem_base + G::TM_Y + half * 32 + ((uint32_t)(lg * 32) << 16);
        const bool tr = SRB_DBG(p.dbg_trace != nullptr) && threadIdx.x == 384;
        const bool no_ld = SRB_DBG(p.dbg_skip & (4 | 128)) != 0, no_st = SRB_DBG(p.dbg_skip & (4 | 256)) != 0;
        const size_t plane = HW;                   // granules per plane
        pdl_wait();                                // x is the previous launch's output; y may still be read by it
        int c = 0, g = g_begin;                    // c: running output-row counter
        int slot = 0; uint32_t yph = 0;            // accumulator ring slot / phase
        RBSeg sg;
        while (rb_next_seg(p, g, g_end, sg, ss_mul, ss_add)) {
            const RBSlot sl = rb_slot_gm(p, sg.ss, grp);
            const int n = sl.n, gx = sl.gx;
            const bool valid = sl.out && gx * 8 + i < p.W;
            // granule index of (n, plane 4*half, row 0, this pixel)
            const size_t px0 = valid ? ((size_t)n * 8 + 4 * half) * plane + (size_t)(gx * 8 + i) : 0;
            const uint4* xg = reinterpret_cast<const uint4*>(p.x) + px0;
            uint4* yg = reinterpret_cast<uint4*>(p.y) + px0;
            uint4 nxt[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                nxt[j] = valid && !no_ld ? rb_ld_res(xg + j * plane + (size_t)sg.y0 * p.W) : make_uint4(0, 0, 0, 0);
#pragma unroll 1
            for (int y = sg.y0; y < sg.y1; ++y, ++c) {
                uint4 rq[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) rq[j] = nxt[j];
                if (tr) rb_trace(p, c, 5);      // output epilogue starts waiting
                mbar_wait<RB_SLEEP>(&yafull[slot], yph, 43);
                if (tr) rb_trace(p, c, 6);      // output accumulator complete
                tc_fence_after();
                uint32_t rr[32];
                tmem_ld32(t_lane + slot * G::CB, rr);
                tmem_ld_wait();
                if (RB_BIAS_MODE == 1) {
                    if (!SRB_DBG(p.dbg_skip & 64)) tmem_zero32(t_lane + slot * G::CB);
                } else if (!SRB_DBG(p.dbg_skip & 64)) {    // re-arm the accumulator row with the bias
                    uint32_t bv[32];
#pragma unroll
                    for (int k = 0; k < 32; ++k) bv[k] = __float_as_uint(s_bias[G::CH + half * 32 + k]);
                    tmem_st32(t_lane + slot * G::CB, bv);
                }
                if constexpr (RB_EARLY_REL) {      // the accumulator row goes back to the convB issuer before the residual add / stores
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) {
                        if constexpr (PAIR) mbar_arrive_cluster(bars_leader + (uint32_t)((48 + slot) * 8));
                        else mbar_arrive(&yaempty[slot]);
                    }
                }
                if (tr) rb_trace(p, c, 10);
                if (y + 1 < sg.y1 && valid && !no_ld) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) nxt[j] = rb_ld_res(xg + j * plane + (size_t)(y + 1) * p.W);
                }
                if (tr) rb_trace(p, c, 8);
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
                            float a = __uint_as_float(rr[8 * j + 2 * k]), b = __uint_as_float(rr[8 * j + 2 * k + 1]);
                            if (RB_BIAS_MODE == 1) { a += p.bias[OFF + 8 * j + 2 * k]; b += p.bias[OFF + 8 * j + 2 * k + 1]; }
                            ov[k] = pack2<BF16>(a + f.x, b + f.y);
                        }
                        if (valid && !no_st) rb_st_out(yg + j * plane + (size_t)y * p.W, make_uint4(ov[0], ov[1], ov[2], ov[3]));
                    }
                };
                if (half == 0) finish(std::integral_constant<int, G::CH>{});
                else finish(std::integral_constant<int, G::CH + 32>{});
                if (tr) rb_trace(p, c, 15);
                if (RB_EARLY_REL && tr) rb_trace(p, c, 9);
                if constexpr (!RB_EARLY_REL) {
                    tmem_st_wait();
                    tc_fence_before();
                    if (tr) rb_trace(p, c, 9);
                    __syncwarp();
                    if (lane == 0) {
                        if constexpr (PAIR) mbar_arrive_cluster(bars_leader + (uint32_t)((48 + slot) * 8));
                        else mbar_arrive(&yaempty[slot]);
                    }
                }
                if (tr) rb_trace(p, c, 7);      // output row done
                if (++slot == G::RY) { slot = 0; yph ^= 1; }
            }
        }
    }

    if (SRB_DBG(p.dbg_trace != nullptr) && threadIdx.x == 0) p.dbg_trace[1024 + 8 * blockIdx.x + 3] = clock64();   // producer done
    tc_fence_before();
    __syncthreads();
    if (SRB_DBG(p.dbg_trace != nullptr) && threadIdx.x == 0) {
        long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        p.dbg_trace[1024 + 8 * blockIdx.x + 4] = clock64();
        p.dbg_trace[1024 + 8 * blockIdx.x + 5] = gt;
        const int launch = (int)p.dbg_trace[1024 + 8 * blockIdx.x + 6];
        if (launch < 400) {
            atomicMin(&p.dbg_trace[2304 + 4 * launch + 2], gt);
            atomicMax(&p.dbg_trace[2304 + 4 * launch + 3], gt);
        }
    }
    if constexpr (PAIR) {
        cluster_sync_all();          // the peer's barriers / TMEM are targets of this CTA's commits and arrivals until both are done
        if (warp == 0) { __syncwarp(); tmem_dealloc_pair(tmem_base, 512); }
    } else if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, 512); }
}

}  // namespace srb
