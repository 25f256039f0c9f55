// Implicit-GEMM convolution on tcgen05/TMEM fed by TMA — the hot kernel of the path.
//
// Replaces, for the reference, every `aten::_convolution` + bias + activation + residual-add group
// reached from utils/models.py:97-98 (Conv), :195-196 (ConvWithoutBN), :208-209 (ResidualBlock1),
// :265-271 (RDB), :316-317 (RRDB), :580-583 (Scaler) and the model.pt graph.
//
// Data layout ("C8 planes"): activations are [N][C/8][H][W][8] fp16/bf16 — a 16-byte granule holds
// 8 channels of one pixel; a plane holds one granule per pixel.  In shared memory a TMA box of this
// tensor is [plane][row][pixel] granules, which IS the canonical K-major SWIZZLE_NONE UMMA operand:
// 8 consecutive pixels of a row are one 8x16B core matrix, rows of the tile are SBO apart and
// channel planes are LBO apart.  A filter tap (dy,dx) is then just a different start address into
// the same halo tile (no swizzle => any 16-byte shift is legal), so one TMA load feeds all taps.
//
// GEMM view: an M-tile is 128 pixels (8 wide x 16 high; DXN: 16 wide x 8 high), N = Cout, K = 16 per
// tcgen05.mma.  Measured on B200 (tools/gpu_ubench.py): an SS-mode M=128 MMA costs
// max(N/2, (4 KB + 32 B*N) / 128 B/clk) cycles and a dependent MMA into the same accumulator cannot
// start for ~85 cycles, so a pipeline step is a SUPER-TILE of NT M-tiles sharing one TMA box whose
// MMAs are issued round-robin into NT independent TMEM accumulators.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (warp-uniform loop, one elected lane issues),
// warps 2..5 = epilogue (TMEM -> regs -> bias / LeakyReLU / scale / residual / pack -> global).
// Two accumulator sets overlap the epilogue of step i with the MMAs of step i+1; a ring of halo-tile
// stages overlaps TMA with MMA; residual granules are prefetched into registers before the
// accumulator is waited for, so the epilogue never sits on a DRAM round trip.
#pragma once
#include <cuda.h>
#include "srb_ptx.cuh"

namespace srb {

enum : int { MODE_STD = 0, MODE_PAIRDX = 1, MODE_DXN = 2, MODE_PAIR4 = 3, MODE_DYF = 4 };
// MODE_DYF (3x3, Cout = 32: the dense blocks' convs and their data-gradient stages — SURVEY a9, a18): the three dy taps are
// FOLDED INTO N.  An M-tile is 8 px x 16 INPUT rows (rows y0-1 .. y0+14 of the halo box); one MMA per (dx, channel pair)
// multiplies it by B = [W(dy=0) | W(dy=1) | W(dy=2)] (N = 96), so accumulator lane (row my, px) holds in column slot s the
// contribution of input row y0-1+my to OUTPUT row y0+my-s.  The epilogue sums out(row my-1) = slot0[lane-8] + slot1[lane] +
// slot2[lane+8] (warp shuffles; the rows at a warp's edges through a shared-memory exchange): 14 output rows per tile.
// Why: an SS-mode MMA re-reads its 4 KB A tile for every K = 16 step, so N = 32 MMAs cannot run faster than 40 clk for 16 clk
// of math (59 clk with the two accumulators of a small-problem super-tile); one N = 96 MMA does three taps in 56-76 clk.
// MODE_PAIR4 (9x9 head, Cin = 3): the input plane holds PIXEL-PAIR granules — granule g of a row = [pixel g: c0 c1 c2 0 |
// pixel g+1: c0 c1 c2 0], stored for g = -1 .. W-1 (row pitch W+1 granules; u8nchw_to_pair4_kernel) — so one 16-byte
// granule carries two dx taps and the two K granules of an MMA (g and g+2, LBO = 32 B) carry four: 9 dy x 3 K-steps = 27
// MMAs per M-tile instead of PAIRDX's 45.

// epilogue flags (mirrored in include/srb200.h)
enum : int {
    F_LRELU = 1,      // LeakyReLU(slope) after bias
    F_RESIDUAL = 2,   // y = out_scale * act(conv + b) + res_scale * residual
    F_TANH = 4,       // tanh after bias (DXN tail conv)
    F_U8OUT = 8,      // DXN: round((t+1)/2*255) -> uint8 NCHW (TanhToArrayImage, utils/models.py:448-451)
    F_PIXSHUF2 = 16,  // fold nn.PixelShuffle(2) into the store addressing (Scaler, utils/models.py:580-583)
    F_F32OUT = 32,    // DXN: fp32 NCHW output
    F_MASK = 64,      // backward: multiply output planes [mask_p0, mask_p1) by LeakyReLU'(saved activation)
    F_HWC_BGR = 128,  // with F_U8OUT: store NHWC with the channels reversed (RGB2BGR + permute of rs.py:65-74, utils/datasets.py:113-124)
};

struct ConvParams {
    int N, H, W;
    int tiles_x, tiles_y, num_tiles;   // super-tiles
    int x_c8_off;            // first input plane
    const uint8_t* wpack;    // packed B operand: ksteps x [2][NOUT][8] halves
    const float* bias;       // NOUT floats (zero padded)
    void* y;                 // C8-plane output (STD/PAIRDX) or NCHW u8/f32 (DXN)
    int y_c8_total, y_c8_off;
    const void* res;         // C8-plane residual, same H, W
    int res_c8_total, res_c8_off;
    const void* res2;        // optional second residual (RRDB outer skip), C8 planes, read with plain loads
    int res2_c8_total, res2_c8_off;
    float res_scale, res2_scale, out_scale, slope;
    int res_nplanes;         // residual is added to output planes < res_nplanes only (dgrad into a dense-block gradient)
    const void* mask;        // F_MASK: saved forward activation (C8 planes, same H, W); sign selects 1 or mask_slope
    int mask_c8_total, mask_c8_off, mask_p0, mask_p1;
    float mask_slope;
    int kblocks;             // WS kernels: number of K blocks (C8IN planes each) streamed per super-tile
    int nmma;                // chain layers: N of this layer's MMAs (a chain compiled for NOUT = 64 also runs 32-output layers); 0 = the kernel's
    int flags;
    int cout;                // real output channels (DXN: 3)
    int dbg_swap;            // debug: swap LBO/SBO of the descriptors (diagnostics only)
    long long* dbg_trace;    // optional per-step clock trace of CTA 0 (tools/gpu_trace.py): [64 steps][8 stamps]
    int dbg_skip;            // bottleneck probes: 1 = no TMA tile loads, 2 = no MMAs, 4 = no global I/O in epilogue
};

// NT M-tiles per super-tile: STD tiles sit side by side in x; PAIRDX / DXN tiles are stacked in y.
// WS (weight streaming): C8IN is the planes of ONE K block; every pipeline stage carries the block's halo box
// and its slice of the filter, so Cin is unbounded (dense-block convs, 96..192 channels).
// KACC: accumulators per M-tile.  2 = the K-steps alternate between two TMEM accumulators that the epilogue adds up: an MMA into
// an accumulator cannot start until the previous one into it has finished (~85 clk), so the two M-tiles of a small-problem
// super-tile give N = 32 MMAs only two independent chains (50 clk per MMA measured in the kernel); four chains reach the
// shared-memory bound (40 clk: the 4 KB A tile is re-read for every MMA).
template <int C8IN, int NOUT, int KH, int KW, int MODE, int NT, bool RES = false, bool WS = false, int KACC = 1>
struct ConvGeom {
    static constexpr bool DYF = MODE == MODE_DYF;
    static constexpr bool XT = MODE == MODE_STD || DYF;           // the NT M-tiles of a super-tile sit side by side in x
    static constexpr int TW = 8;                                  // output tile width
    static constexpr int TH = (MODE == MODE_DXN) ? 8 : 16;        // tile height (DYF: 16 input rows -> 14 output rows)
    static constexpr int SUPER_W = XT ? TW * NT : TW;
    static constexpr int SUPER_H = DYF ? TH - 2 : (MODE == MODE_STD) ? TH : TH * NT;
    static constexpr int PADX = KW / 2, PADY = KH / 2;
    static constexpr int NMMA = DYF ? 3 * NOUT : NOUT;            // N of one MMA = TMEM columns of one M-tile's accumulator
    static constexpr int HALO_W = (MODE == MODE_DXN) ? 16 : (MODE == MODE_PAIR4) ? (SUPER_W + KW - 1 + 2)
                                                         : (SUPER_W + KW - 1 + ((MODE == MODE_PAIRDX) ? (KW & 1) : 0));
    static constexpr int HALO_H = SUPER_H + KH - 1;
    static constexpr int ROWP = HALO_W * 16;                      // bytes between tile rows
    static constexpr int PLANE = HALO_H * ROWP;                   // bytes between channel planes
    static constexpr int TOFF = XT ? TW * 16 : TH * ROWP;         // A-operand offset of M-tile t
    static constexpr int STAGE_BYTES = C8IN * PLANE;              // one TMA box
    static constexpr int A_STRIDE = (STAGE_BYTES + 1023) / 1024 * 1024;
    static constexpr int KSTEPS = (MODE == MODE_STD)      ? KH * KW * (C8IN / 2)
                                  : DYF                   ? KW * (C8IN / 2)
                                  : (MODE == MODE_PAIRDX) ? KH * ((KW + 1) / 2)
                                  : (MODE == MODE_PAIR4)  ? KH * ((KW + 3) / 4)
                                                          : KH * (C8IN / 2);
    static constexpr int WSTEP_BYTES = 2 * NMMA * 16;             // one K=16 slice of B
    static constexpr int WBLK_BYTES = KSTEPS * WSTEP_BYTES;      // filter slice of one K block
    static constexpr int WBYTES = WS ? 0 : WBLK_BYTES;           // resident filter (non-WS)
    static constexpr int STAGE_STRIDE = A_STRIDE + (WS ? (WBLK_BYTES + 1023) / 1024 * 1024 : 0);
    static constexpr int ACC_COLS = 2 * NT * NMMA * KACC;         // two accumulator sets
    static_assert(KACC == 1 || (MODE == MODE_STD && KACC == 2), "K-split accumulators: nine-tap 3x3 kernels");
    static constexpr int TMEM_COLS = (ACC_COLS <= 32) ? 32 : (ACC_COLS <= 64) ? 64 : (ACC_COLS <= 128) ? 128
                                     : (ACC_COLS <= 256) ? 256 : 512;
    // DYF: behind the barriers and the chain area, per epilogue set two buffers of [4 warps][2 directions][8 float4 groups][8 px]
    // for the rows exchanged across warps
    static constexpr int XCH_BUF = 4 * 2 * 8 * 8 * 16;            // 8 KB
    static constexpr int XCH_BYTES = DYF ? 4096 + 2 * 2 * XCH_BUF : 0;      // (4096: the chain area, CHAIN_SMEM_EXTRA below)
    static constexpr int SMEM_FIXED = 1024 /*align slack*/ + (WBYTES + 1023) / 1024 * 1024 + NOUT * 4 + 256 /*barriers*/ + XCH_BYTES;
    static constexpr int max_stages() {
        int s = (227 * 1024 - SMEM_FIXED) / STAGE_STRIDE;
        return s > 4 ? 4 : s;
    }
    static constexpr int STAGES = RES ? 2 : max_stages();
    // RES: the residual M-tiles (8 px x 16 rows x NOUT/8 planes) are TMA-loaded into a ring of smem slots
    static constexpr int RES_TILE_BYTES = (NOUT / 8) * 128 * 16;
    static constexpr int max_rslots() {
        int r = (227 * 1024 - SMEM_FIXED - 2 * STAGE_STRIDE) / RES_TILE_BYTES;
        return r > 6 ? 6 : r;
    }
    static constexpr int RSLOTS = RES ? max_rslots() : 0;
    static constexpr int SMEM_BYTES = SMEM_FIXED + STAGES * STAGE_STRIDE + RSLOTS * RES_TILE_BYTES;
    static_assert(!RES || (RSLOTS >= 2 && MODE == MODE_STD), "residual ring needs two slots (STD mode only)");
    static_assert(!DYF || (KH == 3 && KW == 3 && NOUT == 32 && !RES), "DYF folds the dy taps of a 3x3 conv with 32 outputs");
    static_assert((MODE != MODE_PAIRDX && MODE != MODE_PAIR4) || C8IN == 1, "PAIRDX / PAIR4 pair neighbouring pixels of a single plane");
    static_assert(MODE == MODE_PAIRDX || MODE == MODE_PAIR4 || (C8IN % 2) == 0, "K=16 needs two channel planes per MMA");
    static_assert(MODE != MODE_DXN || (KW == 9 && NOUT == 32), "DXN folds 9 dx taps x 3 couts into N=32");
    static_assert(NOUT % 32 == 0 && NOUT <= 256, "epilogue reads TMEM 32 columns at a time");
    static_assert(ACC_COLS <= 512, "TMEM holds 512 columns");
    static_assert(HALO_W * 4 <= 256 && HALO_H <= 256, "TMA box limits (tensor map is typed as 32-bit words: 4 per granule)");
    static_assert(STAGES >= 2, "needs two halo stages in shared memory");
    static_assert(!WS || XT, "weight streaming is implemented for the 3x3 modes");
};

__device__ __forceinline__ float lrelu(float v, float s) { return v > 0.f ? v : v * s; }
// trace stamp k of step `it` (CTA 0, first 64 steps); costs one predicated store when enabled, a test otherwise
__device__ __forceinline__ void trace(const ConvParams& p, int it, int k) {
    if (SRB_DBG(p.dbg_trace != nullptr) && blockIdx.x == 0 && it < 64) p.dbg_trace[it * 8 + k] = clock64();
}

template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    if constexpr (BF16) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    } else {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
}
template <bool BF16>
__device__ __forceinline__ float2 unpack2(uint32_t u) {
    if constexpr (BF16) {
        return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u));
    } else {
        return __half22float2(*reinterpret_cast<__half2*>(&u));
    }
}
__device__ __forceinline__ uint4 ldg_nc16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// warps: 0 halo producer, 1 MMA issuer, 2 residual producer, 3 spare, 4.. epilogue (EPI_SETS x 4 warps)
constexpr int EPI_SETS = 2;

// First super-tile of this CTA in layer l.  A chain rotates the round-robin assignment from layer to layer: with T tiles on G
// CTAs the T mod G CTAs that get one tile more are different ones in every layer, so no CTA carries the surplus of all layers
// (256 tiles on 148 CTAs, four layers: 7 pipeline steps on the busiest CTA instead of 8).
template <bool CHAIN>
__device__ __forceinline__ int chain_first_tile(int l, int num_tiles) {
    if constexpr (!CHAIN) return (int)blockIdx.x;
    const int g = (int)gridDim.x, extra = num_tiles % g;
    return (int)((blockIdx.x + (unsigned)(l * (g - extra))) % (unsigned)g);
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
// bounded spin until the arrival counter at `f` reaches `target` (acquire at GPU scope)
__device__ __forceinline__ void chain_wait(const unsigned int* f, unsigned int target) {
    long long t0 = 0;
    for (int spin = 0;; ++spin) {
        unsigned int v;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if ((int)(v - target) >= 0) return;
        if ((spin & 63) == 63) {
            if (g_srb_abort) return;
            if (t0 == 0) t0 = clock64();
            else if (clock64() - t0 > (1ll << SRB_WAIT_BUDGET_LOG2)) {
                if (atomicExch((unsigned int*)&g_srb_abort, 1u) == 0u) {
                    g_srb_timeout_info[0] = 0xdead0000u | 60u; g_srb_timeout_info[1] = blockIdx.x; g_srb_timeout_info[2] = threadIdx.x; g_srb_timeout_info[3] = target;
                    unsigned int* h = g_srb_abort_host;
                    if (h != nullptr) { h[1] = 0xdead0000u | 60u; h[2] = blockIdx.x; h[3] = threadIdx.x; h[4] = target; __threadfence_system(); *reinterpret_cast<volatile unsigned int*>(h) = 1u; }
                }
                return;
            }
        }
        __nanosleep(100);
    }
}
constexpr int CONV_THREADS = 128 + 128 * EPI_SETS;

__device__ __forceinline__ float4 lds_f4(uint32_t a) {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(a));
    return r;
}
__device__ __forceinline__ void sts_u4x(uint32_t a, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 lds_u4(uint32_t a) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
    return r;
}

// BWD: compile the backward-only epilogue terms (partial residual, LeakyReLU' mask); the inference kernels are
// epilogue-bound in places and must not carry them.
// ---- CHAIN: several convs of one dense block as ONE launch (conv_chain_kernel) -------------------------------------------
// RDB.forward (utils/models.py:265-271) is a chain o_k = act(conv_k(cat[x, o_0 .. o_{k-1}])) over one dense buffer, and so is
// its backward (training.py).  At a training batch (16 x 64 x 64: 3.5 M-tiles per SM) a launch per conv is mostly latency:
// prologue, first halo, the last epilogue, the wait for the slowest CTA — 12-17 us for 3-9 us of MMA issue.  In a chain the
// persistent CTAs run layer after layer; a super-tile of layer l starts as soon as the 3 x 3 neighbourhood of super-tiles
// of layer l-1 has been stored (per-tile arrival counters in global memory; the neighbourhood's own inputs are then complete
// as well), so neither a grid-wide barrier nor a launch boundary separates the layers.  All CTAs are resident (grid <= #SMs,
// one CTA per SM) and every CTA visits (layer, tile) in the same order, so the waits cannot deadlock; they are bounded anyway.
constexpr int CHAIN_MAX_LAYERS = 8;
constexpr int CHAIN_SMEM_EXTRA = 4096;      // the layers' parameters and biases, behind the barriers
struct ChainCtl {
    const ConvParams* layers;      // [nlayers] device
    const CUtensorMap* tmaps;      // [nlayers] device (64-byte aligned)
    int nlayers;
    unsigned int* flags;           // [nlayers][num_tiles] arrival counters: + EPI warps per finished super-tile, never reset
    unsigned int* epoch;           // launches of this chain completed so far (a tile of this launch is done at (epoch + 1) * 8)
    unsigned int* ticket;          // CTAs of the running launch that have finished
};

template <int C8IN, int NOUT, int KH, int KW, int MODE, int NT, bool RES, bool WS, bool BF16, bool BWD, bool CHAIN, int KACC = 1>
__device__ __forceinline__ void conv_body(const CUtensorMap* tmap_p, const CUtensorMap* tmap_res_p, const ConvParams& p0, const ChainCtl& cc) {
    using G = ConvGeom<C8IN, NOUT, KH, KW, MODE, NT, RES, WS, KACC>;
    const CUtensorMap& tmap = *tmap_p;
    const CUtensorMap& tmap_res = *tmap_res_p;
    const int nl = CHAIN ? cc.nlayers : 1;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* s_w = smem;                                       // WBYTES
    uint8_t* s_a = smem + (G::WBYTES + 1023) / 1024 * 1024;    // STAGES x STAGE_STRIDE
    uint8_t* s_res = s_a + G::STAGES * G::STAGE_STRIDE;        // RSLOTS x RES_TILE_BYTES
    float* s_bias = reinterpret_cast<float*>(s_res + G::RSLOTS * G::RES_TILE_BYTES);
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_bias + NOUT);
    uint64_t* full = bars;                   // [STAGES]
    uint64_t* empty = bars + 4;              // [STAGES]
    uint64_t* tfull = bars + 8;              // [2]
    uint64_t* tempty = bars + 10;            // [2]
    uint64_t* wbar = bars + 12;              // [1]
    uint64_t* rfull = bars + 14;             // [RSLOTS <= 6]
    uint64_t* rempty = bars + 20;            // [RSLOTS <= 6]
    uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 28);
    // CHAIN: the layers' parameters and biases live behind the barriers (the launch adds CHAIN_SMEM_EXTRA bytes)
    ConvParams* s_lay = reinterpret_cast<ConvParams*>(reinterpret_cast<uint8_t*>(bars) + 256);
    float* s_bias_l = reinterpret_cast<float*>(s_lay + CHAIN_MAX_LAYERS);          // [nlayers][NOUT]
    uint8_t* s_xch = reinterpret_cast<uint8_t*>(bars) + 256 + CHAIN_SMEM_EXTRA;                // DYF: rows exchanged between the epilogue warps

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    if (threadIdx.x == 0 && SRB_DBG(p0.dbg_trace != nullptr) && blockIdx.x == 0) p0.dbg_trace[7] = clock64();    // kernel entry

    if constexpr (CHAIN) {
        for (int i = threadIdx.x; i < nl * (int)(sizeof(ConvParams) / 4); i += blockDim.x)
            reinterpret_cast<uint32_t*>(s_lay)[i] = reinterpret_cast<const uint32_t*>(cc.layers)[i];
        for (int i = threadIdx.x; i < nl * NOUT; i += blockDim.x) {
            const ConvParams& q = cc.layers[i / NOUT];              // (a 32-output layer of a NOUT = 64 chain has 32 bias entries)
            s_bias_l[i] = (i % NOUT) < (q.nmma > 0 ? q.nmma : NOUT) ? q.bias[i % NOUT] : 0.f;
        }
    } else {
        for (int i = threadIdx.x; i < NOUT; i += blockDim.x) s_bias[i] = p0.bias[i];
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < G::STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(&tfull[0], 1); mbar_init(&tfull[1], 1);
        mbar_init(&tempty[0], 4 * EPI_SETS); mbar_init(&tempty[1], 4 * EPI_SETS);
        mbar_init(wbar, 1);
        // NT == 1: both epilogue sets work on the same M-tile (alternate 32-column chunks) and both release its residual slot
        for (int s = 0; s < G::RSLOTS; ++s) { mbar_init(&rfull[s], 1); mbar_init(&rempty[s], NT == 1 ? 4 * EPI_SETS : 4); }
        fence_mbar_init();
        tma_prefetch_desc(&tmap);
        if (RES) tma_prefetch_desc(&tmap_res);
    }
    if (warp == 0) { __syncwarp(); tmem_alloc(s_tmem, G::TMEM_COLS); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *s_tmem;
    const ConvParams& p = CHAIN ? s_lay[0] : p0;          // fields common to all layers of a chain: shape, tiling, probes
    if (threadIdx.x == 0 && SRB_DBG(p.dbg_trace != nullptr) && blockIdx.x == 0) p.dbg_trace[15] = clock64();   // prologue done
    // programmatic dependent launch (srb_model_pt_forward chains its launches this way): the prologue above touched only
    // this launch's constants; the next launch may start its own on the SMs this grid vacates, and every warp below that
    // touches activations first waits for the previous launch to complete (a no-op for a normal launch)
    pdl_launch_dependents();

    const int tiles_per_img = p.tiles_x * p.tiles_y;

    if (warp == 0) {
        // ===================== TMA producer (warp-uniform loop, one elected lane issues) =====================
        if (!WS && elect_one()) {
            mbar_arrive_expect_tx(wbar, G::WBYTES);
            for (int off = 0; off < G::WBYTES; off += 16384) {
                int n = G::WBYTES - off < 16384 ? G::WBYTES - off : 16384;
                bulk_load_1d(s_w + off, p.wpack + off, n, wbar);
            }
        }
        __syncwarp();
        pdl_wait();
        int s = 0; uint32_t ph = 0;
        int pit = 0;
        unsigned int dep_target = 0;
        if constexpr (CHAIN) dep_target = (*reinterpret_cast<volatile unsigned int*>(cc.epoch) + 1u) * (4u * EPI_SETS);
#pragma unroll 1
        for (int l = 0; l < nl; ++l) {
        const ConvParams& p = CHAIN ? s_lay[l] : p0;
        const CUtensorMap* tm = CHAIN ? cc.tmaps + l : tmap_p;
        const int kblocks = WS ? p.kblocks : 1;
        const int wblk = (CHAIN && p.nmma > 0) ? G::WBLK_BYTES / G::NMMA * p.nmma : G::WBLK_BYTES;     // this layer's filter slice per K block
        for (int tile = chain_first_tile<CHAIN>(l, p.num_tiles); tile < p.num_tiles; tile += gridDim.x, ++pit) {
            int n = tile / tiles_per_img;
            int r = tile - n * tiles_per_img;
            int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
            if constexpr (CHAIN) {
                if (l > 0) {     // the 3 x 3 neighbourhood of this super-tile in the previous layer must have been stored
                    const int ny = ty + lane / 3 - 1, nx = tx + lane % 3 - 1;
                    if (lane < 9 && ny >= 0 && ny < p.tiles_y && nx >= 0 && nx < p.tiles_x)
                        chain_wait(cc.flags + (size_t)(l - 1) * p.num_tiles + (size_t)n * tiles_per_img + ny * p.tiles_x + nx, dep_target);
                    __syncwarp();
                    fence_proxy_async_global();      // the TMA (async proxy) reads what other SMs stored through the generic proxy
                }
            }
#pragma unroll 1
            for (int kb = 0; kb < kblocks; ++kb) {
                mbar_wait(&empty[s], ph ^ 1, 1);
                if (lane == 0 && kb == 0) trace(p, pit, 0);     // halo load issued
                if (elect_one()) {
                    if SRB_DBG(p.dbg_skip & 1) { mbar_arrive(&full[s]); }
                    else {
                        mbar_arrive_expect_tx(&full[s], G::STAGE_BYTES + (WS ? wblk : 0));
                        tma_load_4d(s_a + s * G::STAGE_STRIDE, tm, &full[s], (tx * G::SUPER_W - G::PADX + (MODE == MODE_PAIR4 ? 1 : 0)) * 4,
                                    ty * G::SUPER_H - G::PADY, p.x_c8_off + kb * C8IN, n);
                        if constexpr (WS) {
                            uint8_t* wdst = s_a + s * G::STAGE_STRIDE + G::A_STRIDE;
                            const uint8_t* wsrc = p.wpack + (size_t)kb * wblk;
                            for (int off = 0; off < wblk; off += 16384) {
                                int nb = wblk - off < 16384 ? wblk - off : 16384;
                                bulk_load_1d(wdst + off, wsrc + off, nb, &full[s]);
                            }
                        }
                    }
                }
                __syncwarp();
                if (++s == G::STAGES) { s = 0; ph ^= 1; }
            }
        }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc0 = umma_idesc_f16(128, G::NMMA, BF16 ? 1 : 0);
        uint32_t a_lbo = (MODE == MODE_PAIRDX) ? 16u : (MODE == MODE_PAIR4) ? 32u : (uint32_t)G::PLANE;
        uint32_t a_sbo = (MODE == MODE_DXN) ? 128u : (uint32_t)G::ROWP;
        uint32_t b_lbo = G::NMMA * 16, b_sbo = 128;
        if (SRB_DBG(p.dbg_swap)) { uint32_t t = a_lbo; a_lbo = a_sbo; a_sbo = t; t = b_lbo; b_lbo = b_sbo; b_sbo = t; }
        // descriptors with a zero start address; the 14-bit (addr >> 4) field is added per MMA
        const uint64_t a_desc0 = umma_desc_nosw(0, a_lbo, a_sbo);
        const uint64_t b_desc0 = umma_desc_nosw(smem_u32(s_w), b_lbo, b_sbo);
        if (!WS) mbar_wait(wbar, 0, 2);
        int s = 0; uint32_t ph = 0; int it = 0;
#pragma unroll 1
        for (int l = 0; l < nl; ++l) {
        const ConvParams& p = CHAIN ? s_lay[l] : p0;
        const int kblocks = WS ? p.kblocks : 1;
        // (chain: N of this layer's MMAs; the accumulators keep the kernel's column stride)
        const bool own_n = CHAIN && p.nmma > 0 && p.nmma != G::NMMA;
        const uint32_t idesc = own_n ? umma_idesc_f16(128, p.nmma, BF16 ? 1 : 0) : idesc0;
        const uint32_t wstep = own_n ? (uint32_t)(2 * p.nmma * 16) : (uint32_t)G::WSTEP_BYTES;
        if (own_n) b_lbo = (uint32_t)p.nmma * 16; else if (CHAIN) b_lbo = G::NMMA * 16;
        for (int tile = chain_first_tile<CHAIN>(l, p.num_tiles); tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int b = it & 1;
            const uint32_t bph = (it >> 1) & 1;
            mbar_wait(&tempty[b], bph ^ 1, 3);
            if (lane == 0) trace(p, it, 1);      // accumulator set free
#pragma unroll 1
            for (int kb = 0; kb < kblocks; ++kb) {
            mbar_wait(&full[s], ph, 4);
            if (lane == 0 && kb == 0) trace(p, it, 2);      // halo landed, MMA issue starts
            tc_fence_after();
            if (elect_one()) {
                const uint64_t a_desc = a_desc0 + (uint64_t)(smem_u32(s_a + s * G::STAGE_STRIDE) >> 4);
                const uint64_t b_desc_s = WS ? umma_desc_nosw(smem_u32(s_a + s * G::STAGE_STRIDE + G::A_STRIDE), b_lbo, b_sbo) : b_desc0;
                const uint32_t d_tmem = tmem_base + b * (NT * G::NMMA * KACC);
                if (!SRB_DBG(p.dbg_skip & 2)) {
                    int ks = 0;
                    if constexpr (G::DYF) {
                        // dy folded into N: the A tile is the 16 input rows themselves (no row shift), dx is a 16-byte shift
#pragma unroll 1
                        for (int dx = 0; dx < KW; ++dx)
#pragma unroll
                            for (int c2 = 0; c2 < C8IN / 2; ++c2, ++ks) {
                                const uint32_t ao = (2 * c2 * G::PLANE + dx * 16) >> 4;
                                const uint64_t bd = b_desc_s + (uint64_t)((ks * wstep) >> 4);
#pragma unroll
                                for (int t = 0; t < NT; ++t)
                                    umma_f16_ss(d_tmem + t * G::NMMA, a_desc + ao + t * (G::TOFF >> 4), bd, idesc, (ks | kb) != 0);
                            }
                    } else if constexpr (MODE == MODE_STD) {
#pragma unroll 1
                        for (int dy = 0; dy < KH; ++dy)
#pragma unroll 1
                            for (int dx = 0; dx < KW; ++dx)
#pragma unroll
                                for (int c2 = 0; c2 < C8IN / 2; ++c2, ++ks) {
                                    const uint32_t ao = (2 * c2 * G::PLANE + dy * G::ROWP + dx * 16) >> 4;
                                    const uint64_t bd = b_desc_s + (uint64_t)((ks * wstep) >> 4);
                                    if constexpr (KACC == 2) {      // K-steps alternate between the tile's two accumulators (c2 = 0 / 1)
#pragma unroll
                                        for (int t = 0; t < NT; ++t)
                                            umma_f16_ss(d_tmem + (t * 2 + (c2 & 1)) * NOUT, a_desc + ao + t * (G::TOFF >> 4), bd, idesc, (ks | kb) >= 2 || kb != 0);
                                    } else {
#pragma unroll
                                    for (int t = 0; t < NT; ++t)
                                        umma_f16_ss(d_tmem + t * NOUT, a_desc + ao + t * (G::TOFF >> 4), bd, idesc, (ks | kb) != 0);
                                    }
                                }
                    } else if constexpr (MODE == MODE_PAIRDX) {
#pragma unroll 1
                        for (int dy = 0; dy < KH; ++dy)
#pragma unroll
                            for (int dxp = 0; dxp < (KW + 1) / 2; ++dxp, ++ks) {
                                const uint32_t ao = (dy * G::ROWP + 2 * dxp * 16) >> 4;
                                const uint64_t bd = b_desc_s + (uint64_t)((ks * wstep) >> 4);
#pragma unroll
                                for (int t = 0; t < NT; ++t)
                                    umma_f16_ss(d_tmem + t * NOUT, a_desc + ao + t * (G::TOFF >> 4), bd, idesc, ks > 0);
                            }
                    } else if constexpr (MODE == MODE_PAIR4) {
#pragma unroll 1
                        for (int dy = 0; dy < KH; ++dy)
#pragma unroll
                            for (int dxq = 0; dxq < (KW + 3) / 4; ++dxq, ++ks) {
                                const uint32_t ao = (dy * G::ROWP + 4 * dxq * 16) >> 4;
                                const uint64_t bd = b_desc_s + (uint64_t)((ks * wstep) >> 4);
#pragma unroll
                                for (int t = 0; t < NT; ++t)
                                    umma_f16_ss(d_tmem + t * NOUT, a_desc + ao + t * (G::TOFF >> 4), bd, idesc, ks > 0);
                            }
                    } else {
#pragma unroll 1
                        for (int dy = 0; dy < KH; ++dy)
#pragma unroll
                            for (int c2 = 0; c2 < C8IN / 2; ++c2, ++ks) {
                                const uint32_t ao = (2 * c2 * G::PLANE + dy * G::ROWP) >> 4;
                                const uint64_t bd = b_desc_s + (uint64_t)((ks * wstep) >> 4);
#pragma unroll
                                for (int t = 0; t < NT; ++t)
                                    umma_f16_ss(d_tmem + t * NOUT, a_desc + ao + t * (G::TOFF >> 4), bd, idesc, ks > 0);
                            }
                    }
                }
                umma_commit(&empty[s]);                          // stage may be refilled once these MMAs have read it
                if (kb == kblocks - 1) umma_commit(&tfull[b]);   // accumulators complete -> epilogue
            }
            __syncwarp();
            if (++s == G::STAGES) { s = 0; ph ^= 1; }
            }
            if (lane == 0) trace(p, it, 3);      // MMA issue done
        }
        }
    } else if (warp == 2) {
        // ===================== residual producer (RES): one TMA box per M-tile into the smem ring =====================
        if constexpr (RES) {
            pdl_wait();
            int rs = 0; uint32_t rph = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
                const int n = tile / tiles_per_img;
                const int r = tile - n * tiles_per_img;
                const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
#pragma unroll 1
                for (int t = 0; t < NT; ++t) {
                    mbar_wait(&rempty[rs], rph ^ 1, 6);
                    if (elect_one()) {
                        if SRB_DBG(p.dbg_skip & 1) { mbar_arrive(&rfull[rs]); }
                        else {
                            mbar_arrive_expect_tx(&rfull[rs], G::RES_TILE_BYTES);
                            tma_load_4d(s_res + rs * G::RES_TILE_BYTES, &tmap_res, &rfull[rs],
                                        (tx * G::SUPER_W + t * G::TW) * 4, ty * G::SUPER_H, p.res_c8_off, n);
                        }
                    }
                    __syncwarp();
                    if (++rs == G::RSLOTS) { rs = 0; rph ^= 1; }
                }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue: EPI_SETS sets of 4 warps; set e handles M-tiles e, e+EPI_SETS, ... =====================
        // (one warp per scheduler cannot hide its own ALU latency; two or more per scheduler can)
        const int lg = warp & 3;                  // TMEM lane group this warp may access
        const int eset = (warp - 4) >> 2;
        pdl_wait();                               // outputs may still be read, residual / mask written, by the previous launch
        const int L = lg * 32 + lane;             // accumulator row = pixel of the M-tile
        const size_t HW = (size_t)p.H * p.W;
        const bool io = !SRB_DBG(p.dbg_skip & 4);
        const uint32_t s_res_u32 = smem_u32(s_res);
        uint32_t xch_par = 0;                      // DYF: which of the set's two exchange buffers the next exchange uses
        int it = 0;
#pragma unroll 1
        for (int l = 0; l < nl; ++l) {
        const ConvParams& p = CHAIN ? s_lay[l] : p0;
        const bool has_res = (p.flags & F_RESIDUAL) != 0;
        const bool do_lrelu = (p.flags & F_LRELU) != 0;
        const uint32_t s_bias_u32 = CHAIN ? smem_u32(s_bias_l + l * NOUT) : smem_u32(s_bias);
        const float* s_bias = CHAIN ? s_bias_l + l * NOUT : reinterpret_cast<const float*>(bars) - NOUT;
        for (int tile = chain_first_tile<CHAIN>(l, p.num_tiles); tile < p.num_tiles; tile += gridDim.x, ++it) {
            const int b = it & 1;
            const uint32_t bph = (it >> 1) & 1;
            const int n = tile / tiles_per_img;
            const int r = tile - n * tiles_per_img;
            const int tyi = r / p.tiles_x, txi = r - tyi * p.tiles_x;
            const uint32_t t_addr0 = tmem_base + b * (NT * G::NMMA * KACC) + ((uint32_t)(lg * 32) << 16);
            if (threadIdx.x == 128) trace(p, it, 4);   // epilogue starts waiting
            // backward of a 32-channel layer: the saved activation granules of this set's first M-tile are requested before
            // the wait for the accumulators (the mask read was an exposed DRAM / L2 round trip per tile)
            constexpr bool MASK_PF = BWD && NOUT <= 64 && G::XT && (CHAIN || NOUT == 32);
            uint4 mpf[MASK_PF ? 4 : 1];
            bool mpf_ok = false;
            if constexpr (MASK_PF) {
                if ((p.flags & F_MASK) && p.mask_p0 == 0 && p.mask_p1 == 4 && eset < NT) {
                    const int y = tyi * G::SUPER_H + (L >> 3) - (G::DYF ? 1 : 0), x = txi * G::SUPER_W + eset * G::TW + (L & 7);
                    if (y < p.H && x < p.W && (!G::DYF || ((L >> 3) >= 1 && (L >> 3) <= 14))) {
                        mpf_ok = true;
                        const uint8_t* mb = reinterpret_cast<const uint8_t*>(p.mask) +
                                            (((size_t)n * p.mask_c8_total + p.mask_c8_off) * HW + (size_t)y * p.W + x) * 16;
#pragma unroll
                        for (int j = 0; j < 4; ++j) mpf[j] = ldg_nc16(mb + (size_t)j * HW * 16);
                    }
                }
            }
            mbar_wait(&tfull[b], bph, 5);
            if (threadIdx.x == 128) trace(p, it, 5);   // accumulators complete
            tc_fence_after();
            if constexpr (MODE != MODE_DXN) {
#pragma unroll 1
                for (int t = 0; t < NT; ++t) {
                    if (NT > 1 && (t % EPI_SETS) != eset) {
                        // not this set's tile — but still observe its residual fill: parity waits are only
                        // sound if no warp ever falls two phases behind a ring slot
                        if constexpr (RES) {
                            const int c = it * NT + t;
                            mbar_wait(&rfull[c % G::RSLOTS], (uint32_t)((c / G::RSLOTS) & 1), 8);
                        }
                        continue;
                    }
                    // (DYF: lane row my holds input row y0-1+my and produces output row y0+my-1; rows 0 and 15 only feed their neighbours)
                    const int y = tyi * G::SUPER_H + (G::XT ? 0 : t * G::TH) + (L >> 3) - (G::DYF ? 1 : 0);
                    const int x = txi * G::SUPER_W + (G::XT ? t * G::TW : 0) + (L & 7);
                    const bool valid = (y < p.H) && (x < p.W) && io && (!G::DYF || ((L >> 3) >= 1 && (L >> 3) <= 14));
                    const size_t pix = (size_t)y * p.W + x;
                    uint32_t rslot_addr = 0;
                    int rslot = 0;
                    if constexpr (RES) {
                        const int c = it * NT + t;
                        rslot = c % G::RSLOTS;
                        mbar_wait(&rfull[rslot], (uint32_t)((c / G::RSLOTS) & 1), 7);
                        rslot_addr = s_res_u32 + rslot * G::RES_TILE_BYTES + L * 16;
                    }
#pragma unroll 1
                    for (int c0 = 0; c0 < NOUT; c0 += 32) {
                        if (NT == 1 && ((c0 >> 5) % EPI_SETS) != eset) continue;   // one M-tile per step: the sets split its columns
                        uint32_t rr[32];
                        if constexpr (G::DYF) {
                            // out(row my-1) = slot0[lane-8] + slot1[lane] + slot2[lane+8]: warp shuffles inside a warp's four rows,
                            // the rows at the warp's edges through shared memory (two buffers alternate: one barrier per exchange)
                            const uint32_t tb = t_addr0 + t * G::NMMA + c0;
                            uint32_t r0[32];
                            tmem_ld32(tb + NOUT, rr);           // slot 1 (dy = 1): this lane's own output row
                            tmem_ld32(tb, r0);                  // slot 0 (dy = 0): belongs to the row below (lane + 8)
                            tmem_ld_wait();
                            const uint32_t xb = smem_u32(s_xch) + (uint32_t)((eset * 2 + (int)xch_par) * G::XCH_BUF);
                            xch_par ^= 1u;
#pragma unroll
                            for (int k = 0; k < 32; ++k) {
                                const float up = __shfl_up_sync(0xffffffffu, __uint_as_float(r0[k]), 8);
                                if (lane >= 8) rr[k] = __float_as_uint(__uint_as_float(rr[k]) + up);
                            }
                            if (lane >= 24) {
#pragma unroll
                                for (int q = 0; q < 8; ++q)
                                    sts_u4x(xb + (uint32_t)((((lg * 2 + 0) * 8 + q) * 8 + (lane - 24)) * 16), make_uint4(r0[4 * q], r0[4 * q + 1], r0[4 * q + 2], r0[4 * q + 3]));
                            }
                            tmem_ld32(tb + 2 * NOUT, r0);       // slot 2 (dy = 2): belongs to the row above (lane - 8)
                            tmem_ld_wait();
#pragma unroll
                            for (int k = 0; k < 32; ++k) {
                                const float dn = __shfl_down_sync(0xffffffffu, __uint_as_float(r0[k]), 8);
                                if (lane < 24) rr[k] = __float_as_uint(__uint_as_float(rr[k]) + dn);
                            }
                            if (lane < 8) {
#pragma unroll
                                for (int q = 0; q < 8; ++q)
                                    sts_u4x(xb + (uint32_t)((((lg * 2 + 1) * 8 + q) * 8 + lane) * 16), make_uint4(r0[4 * q], r0[4 * q + 1], r0[4 * q + 2], r0[4 * q + 3]));
                            }
                            named_bar_sync(1 + eset, 128);
                            if (lane < 8 && lg > 0) {           // slot 0 of the last row of the warp above
#pragma unroll
                                for (int q = 0; q < 8; ++q) {
                                    const float4 f = lds_f4(xb + (uint32_t)(((((lg - 1) * 2 + 0) * 8 + q) * 8 + lane) * 16));
                                    rr[4 * q] = __float_as_uint(__uint_as_float(rr[4 * q]) + f.x); rr[4 * q + 1] = __float_as_uint(__uint_as_float(rr[4 * q + 1]) + f.y);
                                    rr[4 * q + 2] = __float_as_uint(__uint_as_float(rr[4 * q + 2]) + f.z); rr[4 * q + 3] = __float_as_uint(__uint_as_float(rr[4 * q + 3]) + f.w);
                                }
                            }
                            if (lane >= 24 && lg < 3) {         // slot 2 of the first row of the warp below
#pragma unroll
                                for (int q = 0; q < 8; ++q) {
                                    const float4 f = lds_f4(xb + (uint32_t)(((((lg + 1) * 2 + 1) * 8 + q) * 8 + (lane - 24)) * 16));
                                    rr[4 * q] = __float_as_uint(__uint_as_float(rr[4 * q]) + f.x); rr[4 * q + 1] = __float_as_uint(__uint_as_float(rr[4 * q + 1]) + f.y);
                                    rr[4 * q + 2] = __float_as_uint(__uint_as_float(rr[4 * q + 2]) + f.z); rr[4 * q + 3] = __float_as_uint(__uint_as_float(rr[4 * q + 3]) + f.w);
                                }
                            }
                        } else if constexpr (KACC == 2) {
                            uint32_t r1[32];
                            tmem_ld32(t_addr0 + (t * 2) * NOUT + c0, rr);
                            tmem_ld32(t_addr0 + (t * 2 + 1) * NOUT + c0, r1);
                            tmem_ld_wait();
#pragma unroll
                            for (int k = 0; k < 32; ++k) rr[k] = __float_as_uint(__uint_as_float(rr[k]) + __uint_as_float(r1[k]));
                        } else {
                            tmem_ld32(t_addr0 + t * NOUT + c0, rr);
                            tmem_ld_wait();
                        }
                        if (c0 >= p.cout) continue;   // padded output channels
                        if (!(p.flags & F_PIXSHUF2)) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int gj = (c0 >> 3) + j;
                                if (gj * 8 >= p.cout) break;
                                float v[8];
                                const float4 b0 = lds_f4(s_bias_u32 + (c0 + 8 * j) * 4);
                                const float4 b1 = lds_f4(s_bias_u32 + (c0 + 8 * j + 4) * 4);
                                const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                                // (explicitly rounded steps: the same conv must give the same bits in every instantiation — tiled, chained,
                                // with or without the TMA residual ring — and the compiler contracts a * b + c differently in each)
                                for (int i = 0; i < 8; ++i) {
                                    float tt = __fadd_rn(__uint_as_float(rr[8 * j + i]), bb[i]);
                                    if (do_lrelu) tt = fmaxf(tt, __fmul_rn(tt, p.slope));   // LeakyReLU for 0 <= slope <= 1
                                    v[i] = __fmul_rn(tt, p.out_scale);
                                }
                                if (valid) {
                                    if ((RES || has_res) && (!BWD || gj < p.res_nplanes)) {
                                        uint4 q;
                                        if constexpr (RES) q = lds_u4(rslot_addr + gj * 2048);
                                        else q = ldg_nc16(reinterpret_cast<const uint8_t*>(p.res) +
                                                          (((size_t)n * p.res_c8_total + p.res_c8_off + gj) * HW + pix) * 16);
                                        const uint32_t qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                                        for (int i = 0; i < 4; ++i) {
                                            float2 f = unpack2<BF16>(qq[i]);
                                            v[2 * i] = __fmaf_rn(p.res_scale, f.x, v[2 * i]);
                                            v[2 * i + 1] = __fmaf_rn(p.res_scale, f.y, v[2 * i + 1]);
                                        }
                                    }
                                    if (p.res2 != nullptr) {
                                        const uint4 q = ldg_nc16(reinterpret_cast<const uint8_t*>(p.res2) +
                                                                 (((size_t)n * p.res2_c8_total + p.res2_c8_off + gj) * HW + pix) * 16);
                                        const uint32_t qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                                        for (int i = 0; i < 4; ++i) {
                                            float2 f = unpack2<BF16>(qq[i]);
                                            v[2 * i] = __fmaf_rn(p.res2_scale, f.x, v[2 * i]);
                                            v[2 * i + 1] = __fmaf_rn(p.res2_scale, f.y, v[2 * i + 1]);
                                        }
                                    }
                                    if (BWD && (p.flags & F_MASK) && gj >= p.mask_p0 && gj < p.mask_p1) {
                                        // LeakyReLU backward: the in-place activation kept its sign (slope > 0)
                                        uint4 q;
                                        if (MASK_PF && mpf_ok && t == eset) q = mpf[MASK_PF ? j : 0];
                                        else q = ldg_nc16(reinterpret_cast<const uint8_t*>(p.mask) +
                                                          (((size_t)n * p.mask_c8_total + p.mask_c8_off + gj) * HW + pix) * 16);
                                        const uint32_t qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                                        for (int i = 0; i < 4; ++i) {
                                            float2 f = unpack2<BF16>(qq[i]);
                                            if (!(f.x > 0.f)) v[2 * i] = __fmul_rn(v[2 * i], p.mask_slope);
                                            if (!(f.y > 0.f)) v[2 * i + 1] = __fmul_rn(v[2 * i + 1], p.mask_slope);
                                        }
                                    }
                                    uint4 o;
                                    o.x = pack2<BF16>(v[0], v[1]); o.y = pack2<BF16>(v[2], v[3]);
                                    o.z = pack2<BF16>(v[4], v[5]); o.w = pack2<BF16>(v[6], v[7]);
                                    *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.y) +
                                        (((size_t)n * p.y_c8_total + p.y_c8_off + gj) * HW + pix) * 16) = o;
                                }
                            }
                        } else {
                            // PixelShuffle(2): conv channel 32k + 4c + 2i + j  ->  out channel 8k + c at (2y+i, 2x+j)
                            const int k = c0 >> 5;
                            const size_t W2 = (size_t)p.W * 2, HW4 = HW * 4;
#pragma unroll
                            for (int ij = 0; ij < 4; ++ij) {
                                float v[8];
#pragma unroll
                                for (int c = 0; c < 8; ++c) {
                                    float tt = __uint_as_float(rr[4 * c + ij]) + s_bias[c0 + 4 * c + ij];
                                    if (do_lrelu) tt = fmaxf(tt, tt * p.slope);
                                    v[c] = tt * p.out_scale;
                                }
                                if (valid) {
                                    uint4 o;
                                    o.x = pack2<BF16>(v[0], v[1]); o.y = pack2<BF16>(v[2], v[3]);
                                    o.z = pack2<BF16>(v[4], v[5]); o.w = pack2<BF16>(v[6], v[7]);
                                    const size_t opix = (size_t)(2 * y + (ij >> 1)) * W2 + 2 * x + (ij & 1);
                                    *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.y) +
                                        (((size_t)n * p.y_c8_total + p.y_c8_off + k) * HW4 + opix) * 16) = o;
                                }
                            }
                        }
                    }
                    if constexpr (RES) {
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&rempty[rslot]);
                    }
                }
            } else {
                // DXN: column dx*3+co of row (y, x') is the contribution of input column x' to output
                // (y, x' - dx + 4); gather the 9 horizontally shifted partials with warp shuffles.
#pragma unroll 1
                for (int t = eset; t < NT; t += EPI_SETS) {
                    uint32_t rr[32];
                    tmem_ld32(t_addr0 + t * NOUT, rr);
                    tmem_ld_wait();
                    const int y = tyi * G::SUPER_H + t * G::TH + (L >> 4);
                    const int t16 = L & 15;
                    const int x = txi * G::SUPER_W - G::PADX + t16;
                    float acc[3] = {s_bias[0], s_bias[1], s_bias[2]};
#pragma unroll
                    for (int dx = 0; dx < 9; ++dx)
#pragma unroll
                        for (int co = 0; co < 3; ++co)
                            acc[co] += __shfl_sync(0xffffffffu, __uint_as_float(rr[dx * 3 + co]), (lane + dx - 4) & 31);
                    const bool valid = (t16 >= 4) && (t16 < 12) && (y < p.H) && (x < p.W) && io;
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
            tc_fence_before();
            __syncwarp();
            if (threadIdx.x == 128) trace(p, it, 6);       // epilogue done (set 0, lane group 0)
            if (lane == 0) mbar_arrive(&tempty[b]);
            if constexpr (CHAIN) {          // this warp's share of the super-tile is in global memory: count it (release, GPU scope)
                __threadfence();
                __syncwarp();
                if (lane == 0) atomicAdd(cc.flags + (size_t)l * p.num_tiles + tile, 1u);
            }
        }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) { __syncwarp(); tmem_dealloc(tmem_base, G::TMEM_COLS); }
    if (threadIdx.x == 0 && SRB_DBG(p.dbg_trace != nullptr) && blockIdx.x == 0) p.dbg_trace[23] = clock64();   // exit
    if constexpr (CHAIN) {
        // the last CTA to finish closes the launch: every CTA has read the epoch long before (right after griddepcontrol.wait)
        if (threadIdx.x == 0) {
            __threadfence();
            if (atomicAdd(cc.ticket, 1u) == gridDim.x - 1) { *cc.ticket = 0u; __threadfence(); atomicAdd(cc.epoch, 1u); }
        }
    }
}

template <int C8IN, int NOUT, int KH, int KW, int MODE, int NT, bool RES, bool WS, bool BF16, bool BWD = false>
__global__ void __launch_bounds__(CONV_THREADS, 1) conv_tc_kernel(const __grid_constant__ CUtensorMap tmap,
                                                         const __grid_constant__ CUtensorMap tmap_res, const ConvParams p) {
    conv_body<C8IN, NOUT, KH, KW, MODE, NT, RES, WS, BF16, BWD, false>(&tmap, &tmap_res, p, ChainCtl{});
}

// the weight-streaming 3x3 convs of a dense block as one launch (see CHAIN above)
static_assert(CHAIN_MAX_LAYERS * ((int)sizeof(ConvParams) + 64 * 4) <= CHAIN_SMEM_EXTRA, "chain area");
template <int C8IN, int NOUT, int MODE, int NT, bool BF16, bool BWD, int KACC>
__global__ void __launch_bounds__(CONV_THREADS, 1) conv_chain_kernel(const ChainCtl cc) {
    conv_body<C8IN, NOUT, 3, 3, MODE, NT, false, true, BF16, BWD, true, KACC>(cc.tmaps, cc.tmaps, cc.layers[0], cc);
}

}  // namespace srb
