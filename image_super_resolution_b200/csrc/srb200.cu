// libsrb200.so — C ABI (include/srb200.h) over the sm_100a kernels in srb_conv.cuh / srb_elem.cuh.
// Host side: context, weight packing into the UMMA B-operand layout (with BN folding),
// TMA tensor-map construction, kernel dispatch and the model.pt layer schedule.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stddef.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <vector>
#include <algorithm>
#include <map>
#include <tuple>

#include "../../include/srb200.h"
#ifdef SRB_DEBUG
#include "../../include/srb200_debug.h"
#endif
#include "srb_conv.cuh"
#include "srb_elem.cuh"
#ifdef SRB_DEBUG
#include "srb_ubench.cuh"
#endif
#include "srb_wgrad.cuh"
#include "srb_bwd.cuh"
#include "srb_resblock.cuh"
#include "srb_rowconv.cuh"
#include "srb_head.cuh"
#include "srb_c1tail.cuh"
#include "srb_tail.cuh"
#include "srb_optim.cuh"
#include "srb_bn.cuh"

using namespace srb;

// ------------------------------------------------------------------------------------------
// errors
// ------------------------------------------------------------------------------------------
static thread_local char t_err[512] = "";
static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CUDA_TRY(expr)                                                                            \
    do {                                                                                          \
        cudaError_t e__ = (expr);                                                                 \
        if (e__ != cudaSuccess) return fail(SRB_ECUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

extern "C" const char* srb_last_error(void) { return t_err; }
extern "C" int srb_version(void) { return 100; }

// ------------------------------------------------------------------------------------------
// objects
// ------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct srb_ctx {
    int device = 0;
    int num_sms = 0;
    PFN_encodeTiled encode = nullptr;
    // product options (srb_set_option): a cap on the persistent grids and host-side kernel-family selection
    int max_grid = 0;
    int path = 0;             // SRB_PATH_* bits
    unsigned int* h_abort = nullptr;   // host-mapped mirror of the device abort flag (one per device, shared by its contexts)
    // probe state: only the -DSRB_DEBUG build (libsrb200_dbg.so, tools/) ever sets these
    int dbg_swap = 0;
    int dbg_skip = 0;
    bool pdl = false;         // launch the conv kernels with programmatic stream serialization (set by srb_model_pt_forward)
    bool pdl_all = false;     // ... for every conv launch of this context (srb_set_programmatic_launch)
    int dbg_repeat = 0;       // probes: launch the fused block this many extra times per call
    int dbg_threads = 0, dbg_smem = -1;
    long long* dbg_trace = nullptr;
    // pinned + device staging for the *_host entry
    uint8_t* h_pin = nullptr; size_t h_pin_bytes = 0;
    uint8_t* d_io = nullptr; size_t d_io_bytes = 0;
    void* d_ws = nullptr; size_t d_ws_bytes = 0;
    cudaStream_t own_stream = nullptr;
    // optional per-launch timing of srb_model_pt_forward (bench/roofline only)
    int prof = 0;
    std::vector<cudaEvent_t> ev;
    int nev = 0;
};

struct srb_weights {
    int cout = 0, cin = 0, k = 0, dtype = 0;
    int mode = 0, c8in = 0, nout = 0, ksteps = 0;
    int nmma = 0;      // B rows per K granule = N of one MMA: nout, or 3 * nout for the dy-folded 3x3 operands (MODE_DYF)
    int kblocks = 0;   // > 0: weight-streaming layout, blocks of 4 planes (32 channels)
    uint8_t* d_pack = nullptr;
    float* d_bias = nullptr;
    size_t pack_bytes = 0;
};

struct srb_resblock;
struct srb_rowconv;
struct srb_tail9;
struct srb_head;
struct srb_net {
    srb_weights* conv[19];
    srb_resblock* rb[8];     // fused ResidualBlock1 operands (used when the shape qualifies)
    srb_rowconv* rc1 = nullptr;   // conv1 (+ long skip) on the row-fold kernel
    srb_tail9* tail = nullptr;    // conv2 + tanh -> uint8 on the row-fold kernel
    srb_weights* head4 = nullptr; // conv0 packed for pixel-pair input granules (MODE_PAIR4)
    srb_head* head = nullptr;     // Normalize + conv0 straight from the uint8 frame (srb_head.cuh)
    float mean[3], std_[3], maxv, slope;
    int dtype;
};

// The library works on the context's device and leaves the caller's current device as it found it (PyTorch tracks the
// runtime's current device: a replica created for cuda:1 must not redirect the caller's later .cuda() calls).
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int d) { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; if (prev != d) cudaSetDevice(d); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

static unsigned int* s_abort_host[64] = {};

extern "C" int srb_create(int device, srb_ctx** out) {
    if (!out) return fail(SRB_EINVAL, "srb_create: out is NULL");
    int ndev = 0;
    CUDA_TRY(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(SRB_EINVAL, "srb_create: device %d of %d", device, ndev);
    DeviceGuard dev_guard(device);
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(SRB_ENOSUP, "srb_create: device %d is sm_%d%d; this library holds sm_100a kernels only", device,
                    prop.major, prop.minor);
    srb_ctx* c = new srb_ctx();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
        delete c;
        return fail(SRB_ECUDA, "srb_create: cuTensorMapEncodeTiled not available from the driver");
    }
    c->encode = reinterpret_cast<PFN_encodeTiled>(fn);
    if (device < 64 && !s_abort_host[device]) {
        // the kernels' bounded waits (srb_ptx.cuh: mbar_wait) mirror their abort flag here: the host polls it without a sync
        unsigned int* h = nullptr; unsigned int* d = nullptr;
        if (cudaHostAlloc((void**)&h, 32, cudaHostAllocMapped | cudaHostAllocPortable) == cudaSuccess) {
            memset(h, 0, 32);
            if (cudaHostGetDevicePointer((void**)&d, h, 0) == cudaSuccess &&
                cudaMemcpyToSymbol(g_srb_abort_host, &d, sizeof(d)) == cudaSuccess) s_abort_host[device] = h;
            else cudaFreeHost(h);
        }
        cudaGetLastError();
    }
    c->h_abort = device < 64 ? s_abort_host[device] : nullptr;
    *out = c;
    return SRB_OK;
}

extern "C" void srb_destroy(srb_ctx* c) {
    if (!c) return;
    DeviceGuard dev_guard(c->device);
    if (c->h_pin) cudaFreeHost(c->h_pin);
    if (c->d_io) cudaFree(c->d_io);
    if (c->d_ws) cudaFree(c->d_ws);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    for (cudaEvent_t e : c->ev) cudaEventDestroy(e);
    delete c;
}
extern "C" int srb_num_sms(const srb_ctx* c) { return c ? c->num_sms : 0; }
extern "C" int srb_set_programmatic_launch(srb_ctx* c, int on) {
    if (!c) return fail(SRB_EINVAL, "ctx is NULL");
    c->pdl_all = on != 0;
    return SRB_OK;
}
extern "C" int srb_set_option(srb_ctx* c, int key, int value) {
    if (!c) return fail(SRB_EINVAL, "ctx is NULL");
    if (key == SRB_OPT_MAX_GRID) c->max_grid = value > 0 ? value : 0;
    else if (key == SRB_OPT_PATH) c->path = value;
    else return fail(SRB_EINVAL, "srb_set_option: unknown key %d", key);
    return SRB_OK;
}
#ifdef SRB_DEBUG
extern "C" int srb_debug_set(srb_ctx* c, int key, int value) {
    if (!c) return fail(SRB_EINVAL, "ctx is NULL");
    if (key == 0) c->dbg_swap = value;
    else if (key == 1) c->max_grid = value;
    else if (key == 2) { c->dbg_skip = value; c->path = value; }    // (the probe scripts pass kernel-probe and path bits in one word)
    else if (key == 3) c->dbg_repeat = value;
    else if (key == 4) c->dbg_threads = value;
    else if (key == 5) c->dbg_smem = value;
    else return fail(SRB_EINVAL, "srb_debug_set: unknown key %d", key);
    return SRB_OK;
}
#endif

// ------------------------------------------------------------------------------------------
// weight packing: OIHW fp32 (+BN) -> per K-step [2 granules][NOUT rows][8 k] 16-bit (UMMA B, K-major,
// SWIZZLE_NONE: row n of granule g at byte (g*NOUT + n)*16).  The K-step order must match the MMA
// issue loops of conv_tc_kernel.
// ------------------------------------------------------------------------------------------
static uint16_t f2h(float f, int dtype) {
    if (dtype == SRB_DT_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(f); return *reinterpret_cast<uint16_t*>(&h); }
    __half h = __float2half_rn(f);
    return *reinterpret_cast<uint16_t*>(&h);
}

static int round_up(int v, int m) { return (v + m - 1) / m * m; }

// kernel family + packed-operand geometry of a (cout, cin, k) conv
static int conv_geometry(int cout, int cin, int k, int* mode, int* c8in, int* nout, int* ksteps, int* kblocks, bool force_ws = false,
                         bool fold = false) {
    if (k == 9 && cin == 3) { *mode = MODE_PAIRDX; *c8in = 1; *nout = round_up(cout, 32); *ksteps = 9 * 5; }
    else if (k == 9 && cout == 3 && cin % 16 == 0) { *mode = MODE_DXN; *c8in = cin / 8; *nout = 32; *ksteps = 9 * (*c8in / 2); }
    else if (k == 3 && cin % 16 == 0) { *mode = MODE_STD; *c8in = cin / 8; *nout = round_up(cout, 32); *ksteps = 9 * (*c8in / 2); }
    else return fail(SRB_ENOSUP, "srb_pack_conv: no kernel for cout=%d cin=%d k=%d", cout, cin, k);
    if (*nout > 256) return fail(SRB_ENOSUP, "srb_pack_conv: cout=%d > 256", cout);
    *kblocks = (*mode == MODE_STD && (*c8in > 8 || (force_ws && *c8in % 4 == 0))) ? *c8in / 4 : 0;
    if (fold) {      // dy taps folded into N (MODE_DYF): weight-streaming 3x3 convs with 32 outputs
        if (*mode != MODE_STD || !*kblocks || *nout != 32) return fail(SRB_ENOSUP, "srb_alloc_conv_ex: no dy-folded kernel for cout=%d cin=%d k=%d", cout, cin, k);
        *mode = MODE_DYF; *ksteps = 3 * (*c8in / 2);
    }
    if (*kblocks && (*c8in % 4)) return fail(SRB_ENOSUP, "srb_pack_conv: cin=%d must be a multiple of 32 above 64", cin);
    return SRB_OK;
}

extern "C" int srb_alloc_conv(srb_ctx* ctx, int cout, int cin, int k, int dtype, srb_weights** out) {
    return srb_alloc_conv_ex(ctx, cout, cin, k, dtype, 0, out);
}

extern "C" int srb_alloc_conv_ex(srb_ctx* ctx, int cout, int cin, int k, int dtype, int weight_streaming, srb_weights** out) {
    if (!ctx || !out) return fail(SRB_EINVAL, "srb_alloc_conv: NULL argument");
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "srb_alloc_conv: dtype %d", dtype);
    int mode, c8in, nout, ksteps, kblocks;
    if (int rc = conv_geometry(cout, cin, k, &mode, &c8in, &nout, &ksteps, &kblocks, weight_streaming != 0, weight_streaming == 2)) return rc;
    srb_weights* sw = new srb_weights();
    sw->cout = cout; sw->cin = cin; sw->k = k; sw->dtype = dtype;
    sw->mode = mode; sw->c8in = c8in; sw->nout = nout; sw->ksteps = ksteps; sw->kblocks = kblocks;
    sw->nmma = mode == MODE_DYF ? 3 * nout : nout;
    sw->pack_bytes = (size_t)ksteps * 2 * sw->nmma * 8 * 2;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&sw->d_pack, sw->pack_bytes) != cudaSuccess || cudaMalloc(&sw->d_bias, nout * sizeof(float)) != cudaSuccess) {
        srb_free_weights(sw);
        return fail(SRB_ENOMEM, "srb_alloc_conv: cudaMalloc failed");
    }
    CUDA_TRY(cudaMemset(sw->d_pack, 0, sw->pack_bytes));
    CUDA_TRY(cudaMemset(sw->d_bias, 0, nout * sizeof(float)));
    *out = sw;
    return SRB_OK;
}

extern "C" int srb_pack_conv_device(srb_ctx* ctx, srb_weights* w, const float* w_src, int src_o, int src_i, int o_off,
                                    int c_off, int transpose, const float* bias_src, srb_stream_t stream) {
    if (!ctx || !w || !w_src) return fail(SRB_EINVAL, "srb_pack_conv_device: NULL argument");
    const int so_need = transpose ? c_off + w->cin : o_off + w->cout;
    const int si_need = transpose ? o_off + w->cout : c_off + w->cin;
    if (o_off < 0 || c_off < 0 || so_need > src_o || si_need > src_i)
        return fail(SRB_EINVAL, "srb_pack_conv_device: slice (%d+%d, %d+%d) exceeds the source filter [%d][%d]", o_off, w->cout,
                    c_off, w->cin, src_o, src_i);
    PackParams pp;
    pp.mode = w->mode; pp.c8in = w->c8in; pp.nout = w->nout; pp.nmma = w->nmma; pp.kblocks = w->kblocks; pp.cout = w->cout; pp.cin = w->cin;
    pp.k = w->k; pp.src_i = src_i; pp.o_off = o_off; pp.c_off = c_off; pp.transpose = transpose;
    pp.c_lo = 0; pp.c_n = 0; pp.scale = 1.f;
    pp.total = w->pack_bytes / 2;
    cudaStream_t st = (cudaStream_t)stream;
    int grid = (int)((pp.total + 255) / 256);
    if (w->dtype == SRB_DT_BF16) pack_conv_kernel<true><<<grid, 256, 0, st>>>(w_src, (uint16_t*)w->d_pack, pp, transpose ? nullptr : bias_src, w->d_bias);
    else pack_conv_kernel<false><<<grid, 256, 0, st>>>(w_src, (uint16_t*)w->d_pack, pp, transpose ? nullptr : bias_src, w->d_bias);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_pack_conv(srb_ctx* ctx, const float* w, const float* bias, const float* bn4, float bn_eps, int cout,
                             int cin, int k, int dtype, srb_weights** out) {
    if (!ctx || !w || !out) return fail(SRB_EINVAL, "srb_pack_conv: NULL argument");
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "srb_pack_conv: dtype %d", dtype);
    int mode, c8in, nout, ksteps, kblocks;
    if (int rc = conv_geometry(cout, cin, k, &mode, &c8in, &nout, &ksteps, &kblocks)) return rc;

    // BN fold (fuse_conv_and_bn): W' = W * gamma/sqrt(var+eps); b' = (b - mean)*gamma/sqrt(var+eps) + beta
    std::vector<float> scale(cout, 1.f), fb(cout, 0.f);
    for (int o = 0; o < cout; ++o) {
        float b0 = bias ? bias[o] : 0.f;
        if (bn4) {
            float g = bn4[o], be = bn4[cout + o], mu = bn4[2 * cout + o], var = bn4[3 * cout + o];
            float s = g / sqrtf(var + bn_eps);
            scale[o] = s;
            fb[o] = (b0 - mu) * s + be;
        } else fb[o] = b0;
    }
    auto W = [&](int o, int c, int dy, int dx) -> float { return w[(((size_t)o * cin + c) * k + dy) * k + dx] * scale[o]; };

    const size_t step_elems = (size_t)2 * nout * 8;
    std::vector<uint16_t> pack((size_t)ksteps * step_elems, 0);
    std::vector<float> hbias(nout, 0.f);
    int ks = 0;
    if (kblocks) {
        // weight-streaming order: [K block of 32 channels][dy][dx][2 K-steps]
        for (int kb = 0; kb < kblocks; ++kb) for (int dy = 0; dy < 3; ++dy) for (int dx = 0; dx < 3; ++dx) for (int c2 = 0; c2 < 2; ++c2, ++ks)
            for (int g = 0; g < 2; ++g) for (int n = 0; n < cout; ++n) for (int j = 0; j < 8; ++j)
                pack[ks * step_elems + ((size_t)g * nout + n) * 8 + j] = f2h(W(n, (kb * 4 + 2 * c2 + g) * 8 + j, dy, dx), dtype);
        for (int o = 0; o < cout; ++o) hbias[o] = fb[o];
    } else if (mode == MODE_STD) {
        for (int dy = 0; dy < 3; ++dy) for (int dx = 0; dx < 3; ++dx) for (int c2 = 0; c2 < c8in / 2; ++c2, ++ks)
            for (int g = 0; g < 2; ++g) for (int n = 0; n < cout; ++n) for (int j = 0; j < 8; ++j)
                pack[ks * step_elems + ((size_t)g * nout + n) * 8 + j] = f2h(W(n, (2 * c2 + g) * 8 + j, dy, dx), dtype);
        for (int o = 0; o < cout; ++o) hbias[o] = fb[o];
    } else if (mode == MODE_PAIRDX) {
        for (int dy = 0; dy < 9; ++dy) for (int dxp = 0; dxp < 5; ++dxp, ++ks)
            for (int g = 0; g < 2; ++g) {
                int dx = 2 * dxp + g;
                if (dx >= 9) continue;
                for (int n = 0; n < cout; ++n) for (int j = 0; j < cin; ++j)
                    pack[ks * step_elems + ((size_t)g * nout + n) * 8 + j] = f2h(W(n, j, dy, dx), dtype);
            }
        for (int o = 0; o < cout; ++o) hbias[o] = fb[o];
    } else {  // MODE_DXN: row n = dx*3 + co
        for (int dy = 0; dy < 9; ++dy) for (int c2 = 0; c2 < c8in / 2; ++c2, ++ks)
            for (int g = 0; g < 2; ++g) for (int dx = 0; dx < 9; ++dx) for (int co = 0; co < 3; ++co) for (int j = 0; j < 8; ++j)
                pack[ks * step_elems + ((size_t)g * nout + dx * 3 + co) * 8 + j] = f2h(W(co, (2 * c2 + g) * 8 + j, dy, dx), dtype);
        for (int o = 0; o < 3; ++o) hbias[o] = fb[o];
    }

    srb_weights* sw = new srb_weights();
    sw->cout = cout; sw->cin = cin; sw->k = k; sw->dtype = dtype;
    sw->mode = mode; sw->c8in = c8in; sw->nout = nout; sw->nmma = nout; sw->ksteps = ksteps; sw->kblocks = kblocks;
    sw->pack_bytes = pack.size() * 2;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&sw->d_pack, sw->pack_bytes) != cudaSuccess || cudaMalloc(&sw->d_bias, nout * sizeof(float)) != cudaSuccess) {
        delete sw;
        return fail(SRB_ENOMEM, "srb_pack_conv: cudaMalloc failed");
    }
    CUDA_TRY(cudaMemcpy(sw->d_pack, pack.data(), sw->pack_bytes, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(sw->d_bias, hbias.data(), nout * sizeof(float), cudaMemcpyHostToDevice));
    *out = sw;
    return SRB_OK;
}

// 9x9 head conv, Cin = 3, for pixel-pair input granules (MODE_PAIR4, srb_conv.cuh): K-step ks = dy * 3 + dxq; K granule g
// holds taps dx = 4*dxq + 2*g + {0, 1}, 4 k-slots each (3 channels + 0)
extern "C" int srb_pack_conv_pair4(srb_ctx* ctx, const float* w, const float* bias, int cout, int dtype, srb_weights** out) {
    if (!ctx || !w || !out) return fail(SRB_EINVAL, "srb_pack_conv_pair4: NULL argument");
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "srb_pack_conv_pair4: dtype %d", dtype);
    if (cout != 64) return fail(SRB_ENOSUP, "srb_pack_conv_pair4: cout=%d (64 only)", cout);
    const int nout = 64, ksteps = 27;
    const size_t step_elems = (size_t)2 * nout * 8;
    std::vector<uint16_t> pack((size_t)ksteps * step_elems, 0);
    for (int dy = 0; dy < 9; ++dy) for (int dxq = 0; dxq < 3; ++dxq) for (int g = 0; g < 2; ++g) for (int n = 0; n < cout; ++n)
        for (int j = 0; j < 8; ++j) {
            const int dx = 4 * dxq + 2 * g + (j >> 2), c = j & 3;
            if (dx >= 9 || c >= 3) continue;
            pack[(size_t)(dy * 3 + dxq) * step_elems + ((size_t)g * nout + n) * 8 + j] = f2h(w[(((size_t)n * 3 + c) * 9 + dy) * 9 + dx], dtype);
        }
    srb_weights* sw = new srb_weights();
    sw->cout = cout; sw->cin = 3; sw->k = 9; sw->dtype = dtype;
    sw->mode = MODE_PAIR4; sw->c8in = 1; sw->nout = nout; sw->nmma = nout; sw->ksteps = ksteps; sw->kblocks = 0;
    sw->pack_bytes = pack.size() * 2;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&sw->d_pack, sw->pack_bytes) != cudaSuccess || cudaMalloc(&sw->d_bias, nout * sizeof(float)) != cudaSuccess) {
        srb_free_weights(sw);
        return fail(SRB_ENOMEM, "srb_pack_conv_pair4: cudaMalloc failed");
    }
    std::vector<float> hb(nout, 0.f);
    for (int o = 0; o < cout; ++o) hb[o] = bias ? bias[o] : 0.f;
    CUDA_TRY(cudaMemcpy(sw->d_pack, pack.data(), sw->pack_bytes, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(sw->d_bias, hb.data(), nout * sizeof(float), cudaMemcpyHostToDevice));
    *out = sw;
    return SRB_OK;
}

extern "C" void srb_free_weights(srb_weights* w) {
    if (!w) return;
    if (w->d_pack) cudaFree(w->d_pack);
    if (w->d_bias) cudaFree(w->d_bias);
    delete w;
}

// ------------------------------------------------------------------------------------------
// conv dispatch
// ------------------------------------------------------------------------------------------
template <int C8IN, int NOUT, int KH, int KW, int MODE, int NT, bool RES = false, bool WS = false, bool BWD = false>
static int launch_conv(srb_ctx* ctx, const srb_weights* w, const void* x, int x_c8_total, ConvParams p, cudaStream_t st) {
    using G = ConvGeom<C8IN, NOUT, KH, KW, MODE, NT, RES, WS>;
    p.kblocks = WS ? w->kblocks : 1;
    if (w->ksteps != G::KSTEPS * p.kblocks || (size_t)G::WBLK_BYTES * p.kblocks != w->pack_bytes)
        return fail(SRB_EINVAL, "launch_conv: packed weights do not match the kernel geometry");
    p.tiles_x = (p.W + G::SUPER_W - 1) / G::SUPER_W;
    p.tiles_y = (p.H + G::SUPER_H - 1) / G::SUPER_H;
    long long nt = (long long)p.tiles_x * p.tiles_y * p.N;
    if (nt <= 0 || nt > 0x7fffffffLL) return fail(SRB_EINVAL, "launch_conv: %lld tiles", nt);
    p.num_tiles = (int)nt;

    CUtensorMap tmap;
    const cuuint64_t WX = (cuuint64_t)p.W + (MODE == MODE_PAIR4 ? 1 : 0);       // PAIR4: granules -1 .. W-1 are stored
    cuuint64_t gdim[4] = {WX * 4, (cuuint64_t)p.H, (cuuint64_t)x_c8_total, (cuuint64_t)p.N};
    cuuint64_t gstr[3] = {WX * 16, WX * p.H * 16, WX * p.H * 16 * x_c8_total};
    cuuint32_t box[4] = {(cuuint32_t)G::HALO_W * 4, (cuuint32_t)G::HALO_H, (cuuint32_t)C8IN, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    // typed as 32-bit words (4 per 16-byte granule): only zero fill and byte movement matter, and it lifts the
    // 256-element box limit to 64 pixels per row
    CUresult r = ctx->encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32,
                             4, const_cast<void*>(x), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled failed: %d (W=%d H=%d planes=%d N=%d)", (int)r, p.W, p.H, x_c8_total, p.N);

    CUtensorMap tmap_res = tmap;
    if (RES) {
        cuuint64_t rdim[4] = {(cuuint64_t)p.W * 4, (cuuint64_t)p.H, (cuuint64_t)p.res_c8_total, (cuuint64_t)p.N};
        cuuint64_t rstr[3] = {(cuuint64_t)p.W * 16, (cuuint64_t)p.W * p.H * 16, (cuuint64_t)p.W * p.H * 16 * p.res_c8_total};
        cuuint32_t rbox[4] = {(cuuint32_t)G::TW * 4, (cuuint32_t)G::TH, (cuuint32_t)(NOUT / 8), 1};
        r = ctx->encode(&tmap_res, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(p.res), rdim, rstr, rbox, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled (residual) failed: %d", (int)r);
    }
    int grid = p.num_tiles < ctx->num_sms ? p.num_tiles : ctx->num_sms;
    if (ctx->max_grid > 0 && grid > ctx->max_grid) grid = ctx->max_grid;
    p.dbg_swap = ctx->dbg_swap;
    p.dbg_skip = ctx->dbg_skip;
    p.dbg_trace = ctx->dbg_trace;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(CONV_THREADS); cfg.dynamicSmemBytes = G::SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (ctx->pdl || ctx->pdl_all) ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (w->dtype == SRB_DT_BF16) {
        auto kern = conv_tc_kernel<C8IN, NOUT, KH, KW, MODE, NT, RES, WS, true, BWD>;
        CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, tmap, tmap_res, p));
    } else {
        auto kern = conv_tc_kernel<C8IN, NOUT, KH, KW, MODE, NT, RES, WS, false, BWD>;
        CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, tmap, tmap_res, p));
    }
    return SRB_OK;
}

extern "C" int srb_conv2d_fwd(srb_ctx* ctx, const srb_weights* w, const void* x, int x_c8_total, int x_c8_off, void* y,
                              int y_c8_total, int y_c8_off, const void* residual, int res_c8_total, int res_c8_off,
                              float res_scale, float out_scale, float lrelu_slope, int flags, int N, int H, int W,
                              srb_stream_t stream) {
    return srb_conv2d_fwd_res2(ctx, w, x, x_c8_total, x_c8_off, y, y_c8_total, y_c8_off, residual, res_c8_total, res_c8_off,
                               res_scale, nullptr, 0, 0, 0.f, out_scale, lrelu_slope, flags, N, H, W, stream);
}

extern "C" int srb_conv2d_fwd_res2(srb_ctx* ctx, const srb_weights* w, const void* x, int x_c8_total, int x_c8_off, void* y,
                                   int y_c8_total, int y_c8_off, const void* residual, int res_c8_total, int res_c8_off,
                                   float res_scale, const void* res2, int res2_c8_total, int res2_c8_off, float res2_scale,
                                   float out_scale, float lrelu_slope, int flags, int N, int H, int W, srb_stream_t stream) {
    srb_conv_args a;
    memset(&a, 0, sizeof(a));
    a.x = x; a.x_c8_total = x_c8_total; a.x_c8_off = x_c8_off;
    a.y = y; a.y_c8_total = y_c8_total; a.y_c8_off = y_c8_off;
    a.residual = residual; a.res_c8_total = res_c8_total; a.res_c8_off = res_c8_off; a.res_scale = res_scale;
    a.res2 = res2; a.res2_c8_total = res2_c8_total; a.res2_c8_off = res2_c8_off; a.res2_scale = res2_scale;
    a.out_scale = out_scale; a.lrelu_slope = lrelu_slope; a.flags = flags; a.N = N; a.H = H; a.W = W;
    return srb_conv2d_ex(ctx, w, &a, stream);
}

extern "C" int srb_conv2d_ex(srb_ctx* ctx, const srb_weights* w, const srb_conv_args* a, srb_stream_t stream) {
    if (!ctx || !w || !a || !a->x || !a->y) return fail(SRB_EINVAL, "srb_conv2d: NULL argument");
    const int N = a->N, H = a->H, W = a->W, flags = a->flags;
    const void* x = a->x; void* y = a->y; const void* residual = a->residual; const void* res2 = a->res2;
    const int x_c8_total = a->x_c8_total, x_c8_off = a->x_c8_off, y_c8_total = a->y_c8_total, y_c8_off = a->y_c8_off;
    const int res_c8_total = a->res_c8_total, res_c8_off = a->res_c8_off;
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_conv2d_fwd: empty shape %dx%dx%d", N, H, W);
    if (((uintptr_t)x & 15) || ((uintptr_t)y & 15 && w->mode != MODE_DXN) || ((uintptr_t)residual & 15))
        return fail(SRB_EINVAL, "srb_conv2d_fwd: buffers must be 16-byte aligned");
    if (x_c8_off < 0 || x_c8_off + w->c8in > x_c8_total) return fail(SRB_EINVAL, "srb_conv2d_fwd: input plane slice out of range");
    if ((flags & SRB_F_RESIDUAL) && !residual) return fail(SRB_EINVAL, "srb_conv2d_fwd: SRB_F_RESIDUAL without residual");
    if ((long long)H * W * 16 * x_c8_total >= (1ll << 40)) return fail(SRB_EINVAL, "srb_conv2d_fwd: image too large");
    ConvParams p{};
    p.N = N; p.H = H; p.W = W;
    p.x_c8_off = x_c8_off;
    p.wpack = w->d_pack; p.bias = w->d_bias;
    p.y = y; p.y_c8_total = y_c8_total; p.y_c8_off = y_c8_off;
    p.res = residual; p.res_c8_total = res_c8_total; p.res_c8_off = res_c8_off;
    p.res_scale = a->res_scale; p.out_scale = a->out_scale; p.slope = a->lrelu_slope;
    p.res2 = res2; p.res2_c8_total = a->res2_c8_total; p.res2_c8_off = a->res2_c8_off; p.res2_scale = a->res2_scale;
    if (res2 && (((uintptr_t)res2 & 15) || !(flags & SRB_F_RESIDUAL))) return fail(SRB_EINVAL, "srb_conv2d_fwd: res2 needs SRB_F_RESIDUAL and 16-byte alignment");
    p.flags = flags; p.cout = w->cout;
    p.res_nplanes = a->res_nplanes > 0 ? a->res_nplanes : 1 << 30;
    p.mask = a->mask; p.mask_c8_total = a->mask_c8_total; p.mask_c8_off = a->mask_c8_off;
    p.mask_p0 = a->mask_p0; p.mask_p1 = a->mask_p1; p.mask_slope = a->mask_slope;
    if (flags & SRB_F_MASK) {
        if (!a->mask || ((uintptr_t)a->mask & 15)) return fail(SRB_EINVAL, "srb_conv2d: SRB_F_MASK needs a 16-byte aligned mask buffer");
        if (a->mask_p0 < 0 || a->mask_p1 < a->mask_p0 || a->mask_c8_off + a->mask_p1 > a->mask_c8_total)
            return fail(SRB_EINVAL, "srb_conv2d: mask plane range out of bounds");
        if ((w->mode != MODE_STD && w->mode != MODE_DYF) || (flags & SRB_F_PIXSHUF2)) return fail(SRB_ENOSUP, "srb_conv2d: SRB_F_MASK is a 3x3 conv epilogue");
    }
    const bool partial_res = a->res_nplanes > 0 && a->res_nplanes < (w->cout + 7) / 8;
    cudaStream_t st = (cudaStream_t)stream;
    if (w->mode == MODE_DXN) {
        if (!(flags & (SRB_F_U8OUT | SRB_F_F32OUT))) return fail(SRB_EINVAL, "srb_conv2d_fwd: tail conv needs SRB_F_U8OUT or SRB_F_F32OUT");
        if ((flags & SRB_F_HWC_BGR) && !(flags & SRB_F_U8OUT)) return fail(SRB_EINVAL, "srb_conv2d_fwd: SRB_F_HWC_BGR is a uint8 store");
        if (w->c8in == 8) return launch_conv<8, 32, 9, 9, MODE_DXN, 4>(ctx, w, x, x_c8_total, p, st);
        return fail(SRB_ENOSUP, "srb_conv2d_fwd: tail conv with cin=%d", w->cin);
    }
    if (flags & (SRB_F_U8OUT | SRB_F_F32OUT | SRB_F_TANH | SRB_F_HWC_BGR)) return fail(SRB_EINVAL, "srb_conv2d_fwd: tanh/u8/f32/HWC output only on the tail conv");
    int out_planes = (flags & SRB_F_PIXSHUF2) ? w->cout / 32 : (w->cout + 7) / 8;
    if ((flags & SRB_F_PIXSHUF2) && (w->cout % 32)) return fail(SRB_EINVAL, "srb_conv2d_fwd: PixelShuffle(2) needs cout %% 32 == 0");
    if ((flags & SRB_F_PIXSHUF2) && (flags & SRB_F_RESIDUAL)) return fail(SRB_EINVAL, "srb_conv2d_fwd: PIXSHUF2 with residual unsupported");
    const int res_need = partial_res ? a->res_nplanes : out_planes;
    if ((flags & SRB_F_RESIDUAL) && (res_c8_off < 0 || res_c8_off + res_need > res_c8_total)) return fail(SRB_EINVAL, "srb_conv2d_fwd: residual plane slice out of range");
    if (y_c8_off < 0 || y_c8_off + out_planes > y_c8_total) return fail(SRB_EINVAL, "srb_conv2d_fwd: output plane slice out of range");
    if (w->cout % 8) return fail(SRB_ENOSUP, "srb_conv2d_fwd: cout=%d not a multiple of 8", w->cout);
    if (w->mode == MODE_PAIR4) {
        if (x_c8_total != 1 || x_c8_off != 0) return fail(SRB_EINVAL, "srb_conv2d: the pixel-pair head conv reads a single pair-granule plane");
        return launch_conv<1, 64, 9, 9, MODE_PAIR4, 4>(ctx, w, x, x_c8_total, p, st);
    }
    if (w->mode == MODE_PAIRDX) {
        if (w->nout == 64) return launch_conv<1, 64, 9, 9, MODE_PAIRDX, 4>(ctx, w, x, x_c8_total, p, st);
        return fail(SRB_ENOSUP, "srb_conv2d_fwd: head conv with cout=%d", w->cout);
    }
    if (w->mode == MODE_DYF) {      // dy-folded weight-streaming 3x3 conv with 32 outputs, launched on its own (normally a chain layer)
        if ((flags & SRB_F_RESIDUAL) || (flags & SRB_F_PIXSHUF2)) return fail(SRB_ENOSUP, "srb_conv2d: dy-folded convs carry no residual / PixelShuffle");
        const long long mt = (long long)((W + 7) / 8) * ((H + 13) / 14) * N;
        const bool sm_ = mt < (long long)ctx->num_sms * 6 && !(ctx->path & SRB_PATH_LARGE_TILING);
        const bool bw = (flags & SRB_F_MASK) != 0;
        if (bw) return sm_ ? launch_conv<4, 32, 3, 3, MODE_DYF, 2, false, true, true>(ctx, w, x, x_c8_total, p, st)
                           : launch_conv<4, 32, 3, 3, MODE_DYF, 2, false, true, true>(ctx, w, x, x_c8_total, p, st);
        return launch_conv<4, 32, 3, 3, MODE_DYF, 2, false, true, false>(ctx, w, x, x_c8_total, p, st);
    }
    // the TMA residual ring needs the residual on every output plane of a full-width (cout == NOUT) conv
    // Small problems (a training batch: 16 x 64 x 64 px = 512 M-tiles for 148 SMs): with NT = 4 a CTA gets ONE super-tile and
    // its load, MMAs and epilogue run back to back; two-tile super-tiles give every CTA two pipeline steps that overlap.
    const long long mtiles = (long long)((W + 7) / 8) * ((H + 15) / 16) * N;
    const bool small0 = mtiles < (long long)ctx->num_sms * 6 && !(ctx->path & SRB_PATH_LARGE_TILING);   // (debug bit 65536: always the large-problem tiling)
    const bool small = small0 && !((ctx->path & SRB_PATH_WS_LARGE_TILING) && w->kblocks);               // (debug bit 262144: large tiling for the weight-streaming kernels)
#define SRB_STD(C8, NO, NTT, NTS) if (!w->kblocks && w->c8in == C8 && w->nout == NO) { \
        if ((flags & SRB_F_RESIDUAL) && !partial_res && w->cout == NO && res_c8_total >= NO / 8) { \
            if (small) return launch_conv<C8, NO, 3, 3, MODE_STD, NTS, true>(ctx, w, x, x_c8_total, p, st); \
            return launch_conv<C8, NO, 3, 3, MODE_STD, NTT, true>(ctx, w, x, x_c8_total, p, st); } \
        if (small) return launch_conv<C8, NO, 3, 3, MODE_STD, NTS, false>(ctx, w, x, x_c8_total, p, st); \
        return launch_conv<C8, NO, 3, 3, MODE_STD, NTT, false>(ctx, w, x, x_c8_total, p, st); }
    const bool bwd = (flags & SRB_F_MASK) || partial_res || (w->nout > 64 && !w->kblocks && !(flags & SRB_F_PIXSHUF2));
    if (bwd) {   // data-gradient launches of the training step (srb_conv_args: mask / partial residual)
#define SRB_BWD(C8, NO, NTT, NTS) if (!w->kblocks && w->c8in == C8 && w->nout == NO) { \
        if ((flags & SRB_F_RESIDUAL) && !partial_res && w->cout == NO && res_c8_total >= NO / 8) { \
            if (small) return launch_conv<C8, NO, 3, 3, MODE_STD, NTS, true, false, true>(ctx, w, x, x_c8_total, p, st); \
            return launch_conv<C8, NO, 3, 3, MODE_STD, NTT, true, false, true>(ctx, w, x, x_c8_total, p, st); } \
        if (small) return launch_conv<C8, NO, 3, 3, MODE_STD, NTS, false, false, true>(ctx, w, x, x_c8_total, p, st); \
        return launch_conv<C8, NO, 3, 3, MODE_STD, NTT, false, false, true>(ctx, w, x, x_c8_total, p, st); }
        SRB_BWD(4, 64, 4, 2) SRB_BWD(8, 64, 3, 2) SRB_BWD(4, 32, 4, 2) SRB_BWD(8, 32, 4, 2)
        // gathered data gradients of a dense block (training.py): the transposed convs of ALL later layers as one conv over
        // their concatenated output gradients -> the 32 channels of one earlier activation, masked by its LeakyReLU'
        if (w->kblocks && w->nout == 32 && !(flags & SRB_F_RESIDUAL)) {
            if (small) return launch_conv<4, 32, 3, 3, MODE_STD, 2, false, true, true>(ctx, w, x, x_c8_total, p, st);
            return launch_conv<4, 32, 3, 3, MODE_STD, 4, false, true, true>(ctx, w, x, x_c8_total, p, st);
        }
        SRB_BWD(4, 96, 2, 2) SRB_BWD(4, 128, 2, 2) SRB_BWD(4, 160, 1, 1)
        // 64 -> 96: the 108 KB filter leaves no room for a TMA residual ring; residual granules come through plain loads
        if (!w->kblocks && w->c8in == 8 && w->nout == 96) return launch_conv<8, 96, 3, 3, MODE_STD, 2, false, false, true>(ctx, w, x, x_c8_total, p, st);
#undef SRB_BWD
        return fail(SRB_ENOSUP, "srb_conv2d: no backward-epilogue kernel for cin=%d cout=%d", w->cin, w->cout);
    }
    SRB_STD(8, 32, 4, 2) SRB_STD(4, 64, 4, 2) SRB_STD(8, 64, 3, 2) SRB_STD(4, 32, 4, 2)
    if (w->kblocks) {   // dense-block convs (Cin = 96..192): K blocks of 32 channels, filter streamed with the halo
        if (w->nout == 32) {
            if (small) return launch_conv<4, 32, 3, 3, MODE_STD, 2, false, true>(ctx, w, x, x_c8_total, p, st);
            return launch_conv<4, 32, 3, 3, MODE_STD, 4, false, true>(ctx, w, x, x_c8_total, p, st);
        }
        if (w->nout == 64) {
            if ((flags & SRB_F_RESIDUAL) && !partial_res && w->cout == 64) {
                if (small) return launch_conv<4, 64, 3, 3, MODE_STD, 2, true, true>(ctx, w, x, x_c8_total, p, st);
                return launch_conv<4, 64, 3, 3, MODE_STD, 4, true, true>(ctx, w, x, x_c8_total, p, st);
            }
            if (small) return launch_conv<4, 64, 3, 3, MODE_STD, 2, false, true>(ctx, w, x, x_c8_total, p, st);
            return launch_conv<4, 64, 3, 3, MODE_STD, 4, false, true>(ctx, w, x, x_c8_total, p, st);
        }
    }
#undef SRB_STD
    return fail(SRB_ENOSUP, "srb_conv2d_fwd: no 3x3 kernel for cin=%d cout=%d", w->cin, w->cout);
}

// ------------------------------------------------------------------------------------------
// elementwise entry points
// ------------------------------------------------------------------------------------------
static int ew_grid(srb_ctx* ctx, size_t total, int block) {
    size_t b = (total + block - 1) / block;
    size_t cap = (size_t)ctx->num_sms * 8;
    return (int)(b < cap ? (b ? b : 1) : cap);
}

extern "C" int srb_u8nchw_to_c8(srb_ctx* ctx, const uint8_t* in, void* out, const float mean[3], const float std_[3],
                                float maxv, int dtype, int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !in || !out || !mean || !std_) return fail(SRB_EINVAL, "srb_u8nchw_to_c8: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_u8nchw_to_c8: empty shape");
    size_t HW = (size_t)H * W, total = HW * N;
    cudaStream_t st = (cudaStream_t)stream;
    if (HW % 4 == 0 && ((uintptr_t)in & 3) == 0 && HW / 4 < 0x40000000ull) {
        const size_t total4 = total / 4;
        size_t nb = (total4 + 255) / 256, cap = (size_t)ctx->num_sms * 16;
        const int g4 = (int)(nb < cap ? (nb ? nb : 1) : cap);
        if (dtype == SRB_DT_BF16)
            u8nchw_to_c8_x4_kernel<true><<<g4, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, (unsigned)(HW / 4), total4);
        else
            u8nchw_to_c8_x4_kernel<false><<<g4, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, (unsigned)(HW / 4), total4);
        CUDA_TRY(cudaGetLastError());
        return SRB_OK;
    }
    int grid = ew_grid(ctx, total, 256);
    if (dtype == SRB_DT_BF16)
        u8nchw_to_c8_kernel<true><<<grid, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, HW, total);
    else
        u8nchw_to_c8_kernel<false><<<grid, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, HW, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_u8nchw_to_pair4(srb_ctx* ctx, const uint8_t* in, void* out, const float mean[3], const float std_[3],
                                   float maxv, int dtype, int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !in || !out || !mean || !std_) return fail(SRB_EINVAL, "srb_u8nchw_to_pair4: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_u8nchw_to_pair4: empty shape");
    cudaStream_t st = (cudaStream_t)stream;
    if (W % 4 == 0 && ((uintptr_t)in & 3) == 0) {
        const size_t total4 = (size_t)N * H * (W / 4);
        size_t nb = (total4 + 255) / 256, cap = (size_t)ctx->num_sms * 16;
        const int g4 = (int)(nb < cap ? (nb ? nb : 1) : cap);
        if (dtype == SRB_DT_BF16)
            u8nchw_to_pair4_x4_kernel<true><<<g4, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, H, W, total4);
        else
            u8nchw_to_pair4_x4_kernel<false><<<g4, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, H, W, total4);
        CUDA_TRY(cudaGetLastError());
        return SRB_OK;
    }
    const size_t total = (size_t)N * H * ((size_t)W + 1);
    const int grid = ew_grid(ctx, total, 256);
    if (dtype == SRB_DT_BF16)
        u8nchw_to_pair4_kernel<true><<<grid, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, H, W, total);
    else
        u8nchw_to_pair4_kernel<false><<<grid, 256, 0, st>>>(in, (uint4*)out, mean[0], mean[1], mean[2], std_[0], std_[1], std_[2], maxv, H, W, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_f32nchw_to_c8(srb_ctx* ctx, const float* in, void* out, int C, int c8_total, int c8_off, int dtype,
                                 int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !in || !out) return fail(SRB_EINVAL, "srb_f32nchw_to_c8: NULL argument");
    int planes = (C + 7) / 8;
    if (C <= 0 || c8_off < 0 || c8_off + planes > c8_total) return fail(SRB_EINVAL, "srb_f32nchw_to_c8: plane slice out of range");
    size_t HW = (size_t)H * W, total = HW * N * planes;
    int grid = ew_grid(ctx, total, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16) f32nchw_to_c8_kernel<true><<<grid, 256, 0, st>>>(in, (uint4*)out, C, c8_total, c8_off, HW, total);
    else f32nchw_to_c8_kernel<false><<<grid, 256, 0, st>>>(in, (uint4*)out, C, c8_total, c8_off, HW, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_c8_to_f32nchw(srb_ctx* ctx, const void* in, float* out, int C, int c8_total, int c8_off, int dtype,
                                 int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !in || !out) return fail(SRB_EINVAL, "srb_c8_to_f32nchw: NULL argument");
    int planes = (C + 7) / 8;
    if (C <= 0 || c8_off < 0 || c8_off + planes > c8_total) return fail(SRB_EINVAL, "srb_c8_to_f32nchw: plane slice out of range");
    size_t HW = (size_t)H * W, total = HW * N * C;
    int grid = ew_grid(ctx, total, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16) c8_to_f32nchw_kernel<true><<<grid, 256, 0, st>>>(in, out, C, c8_total, c8_off, HW, total);
    else c8_to_f32nchw_kernel<false><<<grid, 256, 0, st>>>(in, out, C, c8_total, c8_off, HW, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// fused optimizer tail over flat fp32 arenas (train.py:58-63)
// ------------------------------------------------------------------------------------------
extern "C" int srb_optim_workspace_floats(void) { return 2 * GSTAT_MAX_BLOCKS + 8; }

extern "C" int srb_grad_stats(srb_ctx* ctx, const float* grads, size_t n, const float* scale_dev, float* stats, float* workspace,
                              srb_stream_t stream) {
    if (!ctx || !grads || !stats || !workspace) return fail(SRB_EINVAL, "srb_grad_stats: NULL argument");
    if (n == 0 || (n & 3)) return fail(SRB_EINVAL, "srb_grad_stats: the arena length must be a positive multiple of 4 floats (pad with zeros)");
    if (((uintptr_t)grads & 15) || ((uintptr_t)workspace & 15)) return fail(SRB_EINVAL, "srb_grad_stats: buffers must be 16-byte aligned");
    const size_t n4 = n / 4;
    size_t b = (n4 + 255) / 256;
    const size_t cap = (size_t)ctx->num_sms * 4 < (size_t)GSTAT_MAX_BLOCKS ? (size_t)ctx->num_sms * 4 : (size_t)GSTAT_MAX_BLOCKS;
    const int grid = (int)(b < cap ? b : cap);
    // workspace: [2 * GSTAT_MAX_BLOCKS] per-block partials, then the arrival counter (zero before the first call; the
    // kernel leaves it at zero)
    unsigned int* counter = reinterpret_cast<unsigned int*>(workspace + 2 * GSTAT_MAX_BLOCKS);
    grad_stats_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(grads), n4, scale_dev, stats, workspace, counter);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_adam_ema_step(srb_ctx* ctx, float* params, const float* grads, float* exp_avg, float* exp_avg_sq, float* ema,
                                 size_t n, const float* scale_dev, float* stats, float* workspace, float max_norm, float lr, float beta1,
                                 float beta2, float eps, float weight_decay, float ema_decay, srb_stream_t stream) {
    if (!ctx || !params || !grads || !exp_avg || !exp_avg_sq || !stats || !workspace) return fail(SRB_EINVAL, "srb_adam_ema_step: NULL argument");
    if (n == 0 || (n & 3)) return fail(SRB_EINVAL, "srb_adam_ema_step: the arena length must be a positive multiple of 4 floats (pad with zeros)");
    if (((uintptr_t)params | (uintptr_t)grads | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq | (uintptr_t)ema) & 15)
        return fail(SRB_EINVAL, "srb_adam_ema_step: arenas must be 16-byte aligned");
    if (!(beta1 >= 0.f && beta1 < 1.f && beta2 >= 0.f && beta2 < 1.f && eps >= 0.f && lr >= 0.f))
        return fail(SRB_EINVAL, "srb_adam_ema_step: lr=%g betas=(%g, %g) eps=%g", lr, beta1, beta2, eps);
    AdamArgs a{};
    a.p = reinterpret_cast<float4*>(params); a.g = reinterpret_cast<const float4*>(grads);
    a.m = reinterpret_cast<float4*>(exp_avg); a.v = reinterpret_cast<float4*>(exp_avg_sq); a.ema = reinterpret_cast<float4*>(ema);
    a.n4 = n / 4; a.scale = scale_dev; a.stats = stats; a.max_norm = max_norm;
    a.counter = reinterpret_cast<unsigned int*>(workspace + 2 * GSTAT_MAX_BLOCKS);
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay; a.ema_decay = ema_decay;
    const int grid = ew_grid(ctx, a.n4, 256);
    adam_ema_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// L1 loss
// ------------------------------------------------------------------------------------------
extern "C" int srb_l1_partials_len(void) { return L1_MAX_BLOCKS; }

static int loss_fwd_impl(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w, float* loss_out,
                         float* partials, srb_stream_t stream, bool sq, const char* who) {
    if (!ctx || !pred || !tgt || !loss_out || !partials) return fail(SRB_EINVAL, "%s: NULL argument", who);
    if (n == 0) return fail(SRB_EINVAL, "%s: empty input (mean of nothing)", who);
    if (((uintptr_t)pred & 15) || ((uintptr_t)tgt & 15)) return fail(SRB_EINVAL, "%s: buffers must be 16-byte aligned", who);
    size_t b = (n / 4 + 255) / 256;
    int grid = (int)(b < (size_t)L1_MAX_BLOCKS ? (b ? b : 1) : L1_MAX_BLOCKS);
    float scale = (float)((double)w / (double)n);
    cudaStream_t st = (cudaStream_t)stream;
#define SRB_LOSS_FWD(DT) \
    if (sq) l1_fwd_kernel<DT, true><<<grid, 256, 0, st>>>(pred, tgt, n, scale, loss_out, partials); \
    else l1_fwd_kernel<DT, false><<<grid, 256, 0, st>>>(pred, tgt, n, scale, loss_out, partials)
    switch (pred_dtype) {
        case 0: SRB_LOSS_FWD(0); break;
        case 1: SRB_LOSS_FWD(1); break;
        case 2: SRB_LOSS_FWD(2); break;
        default: return fail(SRB_EINVAL, "%s: pred_dtype %d", who, pred_dtype);
    }
#undef SRB_LOSS_FWD
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

static int loss_bwd_impl(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w, float grad_scale,
                         const float* grad_out_dev, void* dpred, srb_stream_t stream, bool sq, const char* who) {
    if (!ctx || !pred || !tgt || !dpred) return fail(SRB_EINVAL, "%s: NULL argument", who);
    if (n == 0) return fail(SRB_EINVAL, "%s: empty input", who);
    size_t b = (n + 255) / 256;
    size_t cap = (size_t)ctx->num_sms * 8;
    int grid = (int)(b < cap ? b : cap);
    float g = (float)((double)w * (double)grad_scale / (double)n);
    cudaStream_t st = (cudaStream_t)stream;
#define SRB_LOSS_BWD(DT) \
    if (sq) l1_bwd_kernel<DT, true><<<grid, 256, 0, st>>>(pred, tgt, n, g, dpred, grad_out_dev); \
    else l1_bwd_kernel<DT, false><<<grid, 256, 0, st>>>(pred, tgt, n, g, dpred, grad_out_dev)
    switch (pred_dtype) {
        case 0: SRB_LOSS_BWD(0); break;
        case 1: SRB_LOSS_BWD(1); break;
        case 2: SRB_LOSS_BWD(2); break;
        default: return fail(SRB_EINVAL, "%s: pred_dtype %d", who, pred_dtype);
    }
#undef SRB_LOSS_BWD
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_l1_loss_fwd(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                               float* loss_out, float* partials, srb_stream_t stream) {
    return loss_fwd_impl(ctx, pred, pred_dtype, tgt, n, w, loss_out, partials, stream, false, "srb_l1_loss_fwd");
}

extern "C" int srb_l1_loss_bwd(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                               float grad_scale, void* dpred, srb_stream_t stream) {
    return loss_bwd_impl(ctx, pred, pred_dtype, tgt, n, w, grad_scale, nullptr, dpred, stream, false, "srb_l1_loss_bwd");
}

extern "C" int srb_l1_loss_bwd_dev(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                                   float grad_scale, const float* grad_out_dev, void* dpred, srb_stream_t stream) {
    return loss_bwd_impl(ctx, pred, pred_dtype, tgt, n, w, grad_scale, grad_out_dev, dpred, stream, false, "srb_l1_loss_bwd");
}

extern "C" int srb_mse_loss_fwd(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                                float* loss_out, float* partials, srb_stream_t stream) {
    return loss_fwd_impl(ctx, pred, pred_dtype, tgt, n, w, loss_out, partials, stream, true, "srb_mse_loss_fwd");
}

extern "C" int srb_mse_loss_bwd_dev(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                                    float grad_scale, const float* grad_out_dev, void* dpred, srb_stream_t stream) {
    return loss_bwd_impl(ctx, pred, pred_dtype, tgt, n, w, grad_scale, grad_out_dev, dpred, stream, true, "srb_mse_loss_bwd_dev");
}

// ------------------------------------------------------------------------------------------
// train-mode BatchNorm passes (utils/models.py:75-111 `Conv`; srb_bn.cuh)
// ------------------------------------------------------------------------------------------
static int bn_launch(srb_ctx* ctx, const srb_bn_args* a, int dtype, srb_stream_t stream, int which, const char* who) {
    if (!ctx || !a) return fail(SRB_EINVAL, "%s: NULL argument", who);
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "%s: dtype %d", who, dtype);
    if (a->N <= 0 || a->H <= 0 || a->W <= 0 || a->planes <= 0) return fail(SRB_EINVAL, "%s: empty shape", who);
    if (!a->y || !a->gamma || !a->beta || !a->sums) return fail(SRB_EINVAL, "%s: y / gamma / beta / sums missing", who);
    if (a->y_c8_off < 0 || a->y_c8_off + a->planes > a->y_c8_total) return fail(SRB_EINVAL, "%s: y plane slice out of range", who);
    if (which != 0 && which != 2) {
        if (!a->out) return fail(SRB_EINVAL, "%s: out missing", who);
        if (a->out_c8_off < 0 || a->out_c8_off + a->planes > a->out_c8_total) return fail(SRB_EINVAL, "%s: out plane slice out of range", who);
    }
    if (which == 1) {
        if (a->res && (a->res_c8_off < 0 || a->res_c8_off + a->planes > a->res_c8_total)) return fail(SRB_EINVAL, "%s: res plane slice out of range", who);
        if (a->res2 && (a->res2_c8_off < 0 || a->res2_c8_off + a->planes > a->res2_c8_total)) return fail(SRB_EINVAL, "%s: res2 plane slice out of range", who);
    }
    if (which >= 2) {
        if (!a->da || !a->bsums) return fail(SRB_EINVAL, "%s: da / bsums missing", who);
        if (a->da_c8_off < 0 || a->da_c8_off + a->planes > a->da_c8_total) return fail(SRB_EINVAL, "%s: da plane slice out of range", who);
    }
    if (a->act != 0 && a->act != 1) return fail(SRB_EINVAL, "%s: act %d", who, a->act);
    if ((long long)a->N > 65535) return fail(SRB_EINVAL, "%s: batch %d", who, a->N);
    BNParams p{};
    p.y = a->y; p.y_c8_total = a->y_c8_total; p.y_c8_off = a->y_c8_off;
    p.out = a->out; p.out_c8_total = a->out_c8_total; p.out_c8_off = a->out_c8_off;
    p.planes = a->planes; p.gamma = a->gamma; p.beta = a->beta; p.sums = a->sums;
    p.running_mean = a->running_mean; p.running_var = a->running_var; p.nbt = a->num_batches_tracked;
    if ((p.running_mean == nullptr) != (p.running_var == nullptr)) return fail(SRB_EINVAL, "%s: running_mean and running_var go together", who);
    p.momentum = a->momentum; p.eps = a->eps; p.slope = a->slope; p.act = a->act; p.out_scale = a->out_scale;
    p.res = a->res; p.res_c8_total = a->res_c8_total; p.res_c8_off = a->res_c8_off; p.res_scale = a->res_scale;
    p.res2 = a->res2; p.res2_c8_total = a->res2_c8_total; p.res2_c8_off = a->res2_c8_off; p.res2_scale = a->res2_scale;
    p.da = a->da; p.da_c8_total = a->da_c8_total; p.da_c8_off = a->da_c8_off; p.up_scale = a->up_scale;
    p.bsums = a->bsums; p.dgamma = a->dgamma; p.dbeta = a->dbeta; p.grad_scale = a->grad_scale;
    p.N = a->N; p.H = a->H; p.W = a->W;
    // blocks per (image, plane): enough to fill the GPU a few times over, at least 256 pixels each
    const long long HW = (long long)a->H * a->W;
    long long chunks = (HW + 1023) / 1024;
    const long long want = (long long)ctx->num_sms * 8 / ((long long)a->planes * a->N) + 1;
    if (chunks > want) chunks = want;
    if (chunks < 1) chunks = 1;
    const dim3 grid((unsigned)chunks, (unsigned)a->planes, (unsigned)a->N);
    cudaStream_t st = (cudaStream_t)stream;
    const bool bf = dtype == SRB_DT_BF16;
    switch (which) {
        case 0: if (bf) bn_stats_kernel<true><<<grid, 256, 0, st>>>(p); else bn_stats_kernel<false><<<grid, 256, 0, st>>>(p); break;
        case 1: if (bf) bn_apply_kernel<true><<<grid, 256, 0, st>>>(p); else bn_apply_kernel<false><<<grid, 256, 0, st>>>(p); break;
        case 2: if (bf) bn_bwd_reduce_kernel<true><<<grid, 256, 0, st>>>(p); else bn_bwd_reduce_kernel<false><<<grid, 256, 0, st>>>(p); break;
        default: if (bf) bn_bwd_apply_kernel<true><<<grid, 256, 0, st>>>(p); else bn_bwd_apply_kernel<false><<<grid, 256, 0, st>>>(p); break;
    }
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}
extern "C" int srb_bn_stats(srb_ctx* ctx, const srb_bn_args* a, int dtype, srb_stream_t stream) { return bn_launch(ctx, a, dtype, stream, 0, "srb_bn_stats"); }
extern "C" int srb_bn_apply(srb_ctx* ctx, const srb_bn_args* a, int dtype, srb_stream_t stream) { return bn_launch(ctx, a, dtype, stream, 1, "srb_bn_apply"); }
extern "C" int srb_bn_bwd_reduce(srb_ctx* ctx, const srb_bn_args* a, int dtype, srb_stream_t stream) { return bn_launch(ctx, a, dtype, stream, 2, "srb_bn_bwd_reduce"); }
extern "C" int srb_bn_bwd_apply(srb_ctx* ctx, const srb_bn_args* a, int dtype, srb_stream_t stream) { return bn_launch(ctx, a, dtype, stream, 3, "srb_bn_bwd_apply"); }

// ------------------------------------------------------------------------------------------
// Normalize + conv0 9x9 3 -> 64 + LeakyReLU straight from uint8 frames (srb_head.cuh)
// ------------------------------------------------------------------------------------------
struct srb_head {
    int dtype = 0;
    uint8_t* d_w = nullptr;
    float h_bias[64] = {};
    mutable bool attr_set = false;
};

extern "C" void srb_head_destroy(srb_head* h) {
    if (!h) return;
    if (h->d_w) cudaFree(h->d_w);
    delete h;
}

extern "C" int srb_head_create(srb_ctx* ctx, const float* w, const float* bias, int dtype, srb_head** out) {
    if (!ctx || !w || !bias || !out) return fail(SRB_EINVAL, "srb_head_create: NULL argument");
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "srb_head_create: dtype %d", dtype);
    using G = HDGeom;
    // K-step (dy, kk): K slot k < 15 is pixel offset k / 3, channel k % 3 of the five pixels starting at tap dx = 5 * kk; slot 15 = 0
    std::vector<uint16_t> pk((size_t)G::W_BYTES / 2, 0);
    for (int dy = 0; dy < 9; ++dy) for (int kk = 0; kk < 2; ++kk) for (int k = 0; k < 15; ++k) {
        const int dx = 5 * kk + k / 3, c = k % 3;
        if (dx > 8) continue;
        for (int n = 0; n < 64; ++n)
            pk[(((size_t)(dy * 2 + kk) * 2 + (k >> 3)) * 64 + n) * 8 + (k & 7)] = f2h(w[(((size_t)n * 3 + c) * 9 + dy) * 9 + dx], dtype);
    }
    srb_head* h = new srb_head();
    h->dtype = dtype;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&h->d_w, G::W_BYTES) != cudaSuccess) {
        srb_head_destroy(h);
        return fail(SRB_ENOMEM, "srb_head_create: cudaMalloc failed");
    }
    CUDA_TRY(cudaMemcpy(h->d_w, pk.data(), G::W_BYTES, cudaMemcpyHostToDevice));
    memcpy(h->h_bias, bias, 64 * 4);
    *out = h;
    return SRB_OK;
}

extern "C" int srb_head_fwd(srb_ctx* ctx, const srb_head* h, const uint8_t* in, int in_layout, const float mean[3], const float std_[3],
                            float maxv, void* y, float slope, int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !h || !in || !mean || !std_ || !y) return fail(SRB_EINVAL, "srb_head_fwd: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_head_fwd: empty shape");
    if ((uintptr_t)y & 15) return fail(SRB_EINVAL, "srb_head_fwd: y must be 16-byte aligned");
    if (in_layout != SRB_LAYOUT_NCHW_RGB && in_layout != SRB_LAYOUT_NHWC_BGR && in_layout != SRB_LAYOUT_NHWC_RGB)
        return fail(SRB_EINVAL, "srb_head_fwd: in_layout %d", in_layout);
    using G = HDGeom;
    HeadParams p{};
    p.N = N; p.H = H; p.W = W;
    p.tiles_x = (W + G::TW - 1) / G::TW; p.tiles_y = (H + G::SUPER_H - 1) / G::SUPER_H;
    const long long nt = (long long)N * p.tiles_x * p.tiles_y;
    if (nt > 0x7fffffffLL) return fail(SRB_EINVAL, "srb_head_fwd: too many tiles");
    p.num_tiles = (int)nt;
    p.in = in;
    p.layout = in_layout == SRB_LAYOUT_NCHW_RGB ? HEAD_NCHW_RGB : in_layout == SRB_LAYOUT_NHWC_RGB ? HEAD_NHWC_RGB : HEAD_NHWC_BGR;
    p.fast = (W % 4) == 0 && ((uintptr_t)in & 3) == 0;
    p.wpack = h->d_w;
    memcpy(p.bias, h->h_bias, sizeof(p.bias));
    p.slope = slope;
    for (int i = 0; i < 3; ++i) { p.mean[i] = mean[i]; p.stdv[i] = std_[i]; }
    p.maxv = maxv; p.y = y;
    int grid = p.num_tiles < ctx->num_sms ? p.num_tiles : ctx->num_sms;
    if (ctx->max_grid > 0 && grid > ctx->max_grid) grid = ctx->max_grid;
    if (!h->attr_set) {
        CUDA_TRY(cudaFuncSetAttribute(head_ingest_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(head_ingest_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        h->attr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(G::THREADS); cfg.dynamicSmemBytes = G::SMEM_BYTES;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (ctx->path & SRB_PATH_NO_PDL) ? 0 : 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (h->dtype == SRB_DT_BF16) CUDA_TRY(cudaLaunchKernelEx(&cfg, head_ingest_kernel<true>, p));
    else CUDA_TRY(cudaLaunchKernelEx(&cfg, head_ingest_kernel<false>, p));
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// model.pt graph
// ------------------------------------------------------------------------------------------
extern "C" int srb_model_pt_create(srb_ctx* ctx, const float* const* w, const float* const* b, const float mean[3],
                                   const float std_[3], float maxv, float slope, int dtype, srb_net** out) {
    if (!ctx || !w || !b || !mean || !std_ || !out) return fail(SRB_EINVAL, "srb_model_pt_create: NULL argument");
    srb_net* net = new srb_net();
    memset(net->conv, 0, sizeof(net->conv));
    memset(net->rb, 0, sizeof(net->rb));
    for (int i = 0; i < 3; ++i) { net->mean[i] = mean[i]; net->std_[i] = std_[i]; }
    net->maxv = maxv; net->slope = slope; net->dtype = dtype;
    int rc = SRB_OK;
    for (int i = 0; i < 19 && rc == SRB_OK; ++i) {
        int cout, cin, k;
        if (i == 0) { cout = 64; cin = 3; k = 9; }
        else if (i == 18) { cout = 3; cin = 64; k = 9; }
        else if (i == 17) { cout = 64; cin = 64; k = 3; }
        else if (i & 1) { cout = 32; cin = 64; k = 3; }
        else { cout = 64; cin = 32; k = 3; }
        if (!w[i] || !b[i]) rc = fail(SRB_EINVAL, "srb_model_pt_create: weight %d is NULL", i);
        else rc = srb_pack_conv(ctx, w[i], b[i], nullptr, 0.f, cout, cin, k, dtype, &net->conv[i]);
    }
    for (int i = 0; i < 8 && rc == SRB_OK; ++i)
        rc = srb_resblock_create(ctx, w[1 + 2 * i], b[1 + 2 * i], w[2 + 2 * i], b[2 + 2 * i], dtype, &net->rb[i]);
    if (rc == SRB_OK) rc = srb_rowconv_create(ctx, w[17], b[17], dtype, &net->rc1);
    if (rc == SRB_OK) rc = srb_tail9_create(ctx, w[18], b[18], dtype, &net->tail);
    if (rc == SRB_OK) rc = srb_pack_conv_pair4(ctx, w[0], b[0], 64, dtype, &net->head4);
    if (rc == SRB_OK) rc = srb_head_create(ctx, w[0], b[0], dtype, &net->head);
    if (rc != SRB_OK) { srb_model_pt_destroy(net); return rc; }
    *out = net;
    return SRB_OK;
}

extern "C" void srb_model_pt_destroy(srb_net* net) {
    if (!net) return;
    for (int i = 0; i < 19; ++i) srb_free_weights(net->conv[i]);
    for (int i = 0; i < 8; ++i) srb_resblock_destroy(net->rb[i]);
    srb_rowconv_destroy(net->rc1);
    srb_tail9_destroy(net->tail);
    srb_free_weights(net->head4);
    srb_head_destroy(net->head);
    delete net;
}

static size_t align256(size_t v) { return (v + 255) / 256 * 256; }

extern "C" size_t srb_model_pt_workspace_bytes(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    size_t P = (size_t)N * H * W;
    return align256((size_t)N * H * (W + 1) * 16) + 3 * align256(P * 128) + align256(P * 64);
}

// kernels one forward enqueues: head (ingest + conv0), 8 fused blocks (or 16 convs when the width does not qualify), conv1, tail
extern "C" int srb_model_pt_launches(int H, int W) { return srb_resblock_supported(H, W) ? (srb_c1tail_supported(H, W) ? 10 : 11) : 19; }

#ifdef SRB_DEBUG
// measurement only: per-step clock trace of CTA 0 of the next conv launches (512 int64 device buffer, or NULL)
extern "C" int srb_debug_trace(srb_ctx* ctx, void* dev_buf512) {
    if (!ctx) return fail(SRB_EINVAL, "ctx is NULL");
    ctx->dbg_trace = (long long*)dev_buf512;
    return SRB_OK;
}
#endif

static int report_abort(srb_ctx* ctx, const unsigned int* info) {
    unsigned int zero = 0;
    cudaMemcpyToSymbol(g_srb_abort, &zero, sizeof(zero));
    if (ctx->h_abort) memset(ctx->h_abort, 0, 32);
    return fail(SRB_ECUDA, "kernel pipeline timeout: wait site %u, block %u, thread %u, parity %u — results of the launches "
                "since the last check are invalid", info[0] & 0xffffu, info[1], info[2], info[3]);
}

// Non-blocking: has a kernel of this device reported a pipeline timeout (srb_ptx.cuh: mbar_wait) since the last check?
// Reads the host-mapped mirror of the abort flag — no device synchronisation — so the Python layer calls it at every point
// where it hands results to the caller.  On a timeout the flag is cleared (after a device sync) and SRB_ECUDA is returned.
extern "C" int srb_poll_abort(srb_ctx* ctx) {
    if (!ctx) return fail(SRB_EINVAL, "ctx is NULL");
    if (!ctx->h_abort || !*reinterpret_cast<volatile unsigned int*>(ctx->h_abort)) return SRB_OK;
    DeviceGuard dev_guard(ctx->device);
    cudaDeviceSynchronize();
    unsigned int info[4] = {ctx->h_abort[1], ctx->h_abort[2], ctx->h_abort[3], ctx->h_abort[4]};
    return report_abort(ctx, info);
}

// Synchronise the device and report a kernel-side protocol timeout (see mbar_wait) as an error.
extern "C" int srb_check_device(srb_ctx* ctx) {
    if (!ctx) return fail(SRB_EINVAL, "ctx is NULL");
    DeviceGuard dev_guard(ctx->device);
    CUDA_TRY(cudaDeviceSynchronize());
    unsigned int flag = 0, info[4] = {0, 0, 0, 0};
    CUDA_TRY(cudaMemcpyFromSymbol(&flag, g_srb_abort, sizeof(flag)));
    if (flag) {
        CUDA_TRY(cudaMemcpyFromSymbol(info, g_srb_timeout_info, sizeof(info)));
        return report_abort(ctx, info);
    }
    return SRB_OK;
}

extern "C" int srb_profile_enable(srb_ctx* ctx, int on) {
    if (!ctx) return fail(SRB_EINVAL, "ctx is NULL");
    ctx->prof = on;
    ctx->nev = 0;
    return SRB_OK;
}

// elapsed ms between consecutive events recorded by the last profiled forward; returns the count written
extern "C" int srb_profile_read(srb_ctx* ctx, float* ms, int max_n) {
    if (!ctx || !ms) return fail(SRB_EINVAL, "srb_profile_read: NULL argument");
    if (ctx->nev < 2) return 0;
    if (cudaEventSynchronize(ctx->ev[ctx->nev - 1]) != cudaSuccess) return fail(SRB_ECUDA, "srb_profile_read: event sync failed");
    int n = 0;
    for (int i = 0; i + 1 < ctx->nev && n < max_n; ++i, ++n)
        if (cudaEventElapsedTime(&ms[n], ctx->ev[i], ctx->ev[i + 1]) != cudaSuccess) return fail(SRB_ECUDA, "srb_profile_read: elapsed failed");
    return n;
}

// level 1: an event after every launch (serialises the programmatic overlap: per-kernel durations in isolation);
// level 2: events only around the trunk's eight launches (coarse = true marks): their average duration in the pipeline
static void prof_mark(srb_ctx* ctx, cudaStream_t st, bool first = false, bool coarse = false) {
    if (!ctx->prof) return;
    if (ctx->prof == 2 && !coarse && !first) return;
    if (first) ctx->nev = 0;
    if ((int)ctx->ev.size() <= ctx->nev) {
        cudaEvent_t e;
        cudaEventCreate(&e);
        ctx->ev.push_back(e);
    }
    cudaEventRecord(ctx->ev[ctx->nev++], st);
}

extern "C" int srb_model_pt_forward(srb_ctx* ctx, srb_net* net, const uint8_t* in, uint8_t* out, int N, int H, int W,
                                    void* ws, size_t ws_bytes, srb_stream_t stream) {
    return srb_model_pt_forward_frames(ctx, net, in, out, SRB_LAYOUT_NCHW_RGB, N, H, W, ws, ws_bytes, stream);
}

extern "C" int srb_model_pt_forward_frames(srb_ctx* ctx, srb_net* net, const uint8_t* in, uint8_t* out, int out_layout, int N, int H, int W,
                                    void* ws, size_t ws_bytes, srb_stream_t stream) {
    return srb_model_pt_forward_io(ctx, net, in, SRB_LAYOUT_NCHW_RGB, out, out_layout, N, H, W, ws, ws_bytes, stream);
}

extern "C" int srb_model_pt_forward_io(srb_ctx* ctx, srb_net* net, const uint8_t* in, int in_layout, uint8_t* out, int out_layout, int N, int H,
                                       int W, void* ws, size_t ws_bytes, srb_stream_t stream) {
    if (!ctx || !net || !in || !out || !ws) return fail(SRB_EINVAL, "srb_model_pt_forward: NULL argument");
    if (in_layout != SRB_LAYOUT_NCHW_RGB && in_layout != SRB_LAYOUT_NHWC_BGR && in_layout != SRB_LAYOUT_NHWC_RGB)
        return fail(SRB_EINVAL, "srb_model_pt_forward_io: in_layout %d", in_layout);
    if (in_layout != SRB_LAYOUT_NCHW_RGB && (ctx->path & (SRB_PATH_C8_HEAD | SRB_PATH_PAIR4_HEAD)))
        return fail(SRB_ENOSUP, "srb_model_pt_forward_io: HWC frames are read by the ingest head only (SRB_PATH_C8_HEAD / SRB_PATH_PAIR4_HEAD take NCHW)");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_model_pt_forward: empty shape %dx%dx%d", N, H, W);
    size_t need = srb_model_pt_workspace_bytes(N, H, W);
    if (ws_bytes < need) return fail(SRB_EINVAL, "srb_model_pt_forward: workspace %zu < %zu bytes", ws_bytes, need);
    if ((uintptr_t)ws & 255) return fail(SRB_EINVAL, "srb_model_pt_forward: workspace must be 256-byte aligned");
    if (out_layout != SRB_LAYOUT_NCHW_RGB && out_layout != SRB_LAYOUT_NHWC_BGR) return fail(SRB_EINVAL, "srb_model_pt_forward_frames: out_layout %d", out_layout);
    const int oflags = SRB_F_TANH | SRB_F_U8OUT | (out_layout == SRB_LAYOUT_NHWC_BGR ? SRB_F_HWC_BGR : 0);
    size_t P = (size_t)N * H * W;
    uint8_t* base = (uint8_t*)ws;
    void* in8 = base; base += align256((size_t)N * H * (W + 1) * 16);   // one input plane (pixel-pair granules: W + 1 per row)
    void* x0 = base; base += align256(P * 128);
    void* ra = base; base += align256(P * 128);
    void* rb = base; base += align256(P * 128);
    void* hid = base;
    const float s = net->slope;
    int rc;
    cudaStream_t pst = (cudaStream_t)stream;
    struct PdlScope { srb_ctx* c; ~PdlScope() { c->pdl = false; } } pdl_scope{ctx};
    ctx->pdl = !(ctx->path & SRB_PATH_NO_PDL);
    prof_mark(ctx, pst, true);
    if (!(ctx->path & (SRB_PATH_C8_HEAD | SRB_PATH_PAIR4_HEAD))) {
        // Normalize + conv0 in one launch, reading the uint8 frame itself (18 MMAs per M-tile, no normalised plane in HBM)
        if ((rc = srb_head_fwd(ctx, net->head, in, in_layout, net->mean, net->std_, net->maxv, x0, s, N, H, W, stream))) return rc;
    } else if (!(ctx->path & SRB_PATH_C8_HEAD)) {      // head conv on pixel-pair granules: 27 MMAs per M-tile instead of 45
        if ((rc = srb_u8nchw_to_pair4(ctx, in, in8, net->mean, net->std_, net->maxv, net->dtype, N, H, W, stream))) return rc;
        prof_mark(ctx, pst);
        if ((rc = srb_conv2d_fwd(ctx, net->head4, in8, 1, 0, x0, 8, 0, nullptr, 0, 0, 0.f, 1.f, s, SRB_F_LRELU, N, H, W, stream))) return rc;
    } else {
        if ((rc = srb_u8nchw_to_c8(ctx, in, in8, net->mean, net->std_, net->maxv, net->dtype, N, H, W, stream))) return rc;
        prof_mark(ctx, pst);
        if ((rc = srb_conv2d_fwd(ctx, net->conv[0], in8, 1, 0, x0, 8, 0, nullptr, 0, 0, 0.f, 1.f, s, SRB_F_LRELU, N, H, W, stream))) return rc;
    }
    prof_mark(ctx, pst, false, true);
    void* cur = x0;
    const bool fused = srb_resblock_supported(H, W) && !(ctx->path & SRB_PATH_UNFUSED_TRUNK);
    for (int i = 0; i < 8; ++i) {
        void* nxt = (i & 1) ? rb : ra;
        if (fused) {
            if ((rc = srb_resblock_fwd(ctx, net->rb[i], cur, nxt, s, N, H, W, stream))) return rc;
            prof_mark(ctx, pst, false, i == 7);
            cur = nxt;
            continue;
        }
        if ((rc = srb_conv2d_fwd(ctx, net->conv[1 + 2 * i], cur, 8, 0, hid, 4, 0, nullptr, 0, 0, 0.f, 1.f, s, SRB_F_LRELU, N, H, W, stream))) return rc;
        prof_mark(ctx, pst);
        if ((rc = srb_conv2d_fwd(ctx, net->conv[2 + 2 * i], hid, 4, 0, nxt, 8, 0, cur, 8, 0, 1.f, 1.f, 0.f, SRB_F_RESIDUAL, N, H, W, stream))) return rc;
        prof_mark(ctx, pst);
        cur = nxt;
    }
    // cur == rb after 8 blocks; long skip adds conv0's (post-activation) output
    if (fused && !(ctx->path & (SRB_PATH_TILED_CONV1 | SRB_PATH_TILED_TAIL | SRB_PATH_SPLIT_C1TAIL)) && srb_c1tail_supported(H, W)) {
        // conv1 + skip + conv2 + tanh -> uint8 as one launch: conv1's result never leaves the SM
        if ((rc = srb_c1tail_fwd(ctx, net->rc1, net->tail, cur, x0, out, oflags, N, H, W, stream))) return rc;
        prof_mark(ctx, pst, false, true);
        return SRB_OK;
    }
    if (fused && !(ctx->path & SRB_PATH_TILED_CONV1)) {
        if ((rc = srb_rowconv_fwd(ctx, net->rc1, cur, x0, ra, 0.f, 0, N, H, W, stream))) return rc;
    } else if ((rc = srb_conv2d_fwd(ctx, net->conv[17], cur, 8, 0, ra, 8, 0, x0, 8, 0, 1.f, 1.f, 0.f, SRB_F_RESIDUAL, N, H, W, stream))) return rc;
    prof_mark(ctx, pst);
    if (fused && !(ctx->path & SRB_PATH_TILED_TAIL)) {
        if ((rc = srb_tail9_fwd(ctx, net->tail, ra, out, oflags, N, H, W, stream))) return rc;
    } else if ((rc = srb_conv2d_fwd(ctx, net->conv[18], ra, 8, 0, out, 0, 0, nullptr, 0, 0, 0.f, 1.f, 0.f, oflags, N, H, W, stream))) return rc;
    prof_mark(ctx, pst, false, true);
    return SRB_OK;
}

extern "C" int srb_model_pt_forward_host(srb_ctx* ctx, srb_net* net, const uint8_t* in_host, uint8_t* out_host, int N, int H, int W) {
    if (!ctx || !net || !in_host || !out_host) return fail(SRB_EINVAL, "srb_model_pt_forward_host: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_model_pt_forward_host: empty shape");
    DeviceGuard dev_guard(ctx->device);
    size_t io = (size_t)N * 3 * H * W, io_al = align256(io);
    size_t ws = srb_model_pt_workspace_bytes(N, H, W);
    if (!ctx->own_stream) CUDA_TRY(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
    if (ctx->h_pin_bytes < 2 * io_al) {
        if (ctx->h_pin) cudaFreeHost(ctx->h_pin);
        ctx->h_pin = nullptr; ctx->h_pin_bytes = 0;
        CUDA_TRY(cudaMallocHost((void**)&ctx->h_pin, 2 * io_al));
        ctx->h_pin_bytes = 2 * io_al;
    }
    if (ctx->d_io_bytes < 2 * io_al) {
        if (ctx->d_io) cudaFree(ctx->d_io);
        ctx->d_io = nullptr; ctx->d_io_bytes = 0;
        CUDA_TRY(cudaMalloc((void**)&ctx->d_io, 2 * io_al));
        ctx->d_io_bytes = 2 * io_al;
    }
    if (ctx->d_ws_bytes < ws) {
        if (ctx->d_ws) cudaFree(ctx->d_ws);
        ctx->d_ws = nullptr; ctx->d_ws_bytes = 0;
        CUDA_TRY(cudaMalloc(&ctx->d_ws, ws));
        ctx->d_ws_bytes = ws;
    }
    cudaStream_t st = ctx->own_stream;
    memcpy(ctx->h_pin, in_host, io);
    CUDA_TRY(cudaMemcpyAsync(ctx->d_io, ctx->h_pin, io, cudaMemcpyHostToDevice, st));
    int rc = srb_model_pt_forward(ctx, net, ctx->d_io, ctx->d_io + io_al, N, H, W, ctx->d_ws, ctx->d_ws_bytes, st);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(ctx->h_pin + io_al, ctx->d_io + io_al, io, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if ((rc = srb_poll_abort(ctx))) return rc;
    memcpy(out_host, ctx->h_pin + io_al, io);
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// measurement tooling: tcgen05.mma micro-benchmark (see srb_ubench.cuh) — debug build only
// ------------------------------------------------------------------------------------------
#ifdef SRB_DEBUG
extern "C" int srb_ubench_mma(srb_ctx* ctx, int n, int layout, int racc, int iters, int m64, int grid, long long* out2_host) {
    if (!ctx || !out2_host) return fail(SRB_EINVAL, "srb_ubench_mma: NULL argument");
    if (racc < 1 || racc * n * (m64 > 0 ? m64 : 1) > 512 || m64 > 4) return fail(SRB_EINVAL, "srb_ubench_mma: racc*n*warps > 512 TMEM columns");
    long long* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, 16));
    const int smem = 100 * 1024;
#define UB(NN) case NN: CUDA_TRY(cudaFuncSetAttribute(ubench_mma_kernel<NN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
               ubench_mma_kernel<NN><<<grid, 128, smem>>>(layout, racc, iters, m64, d); break;
    switch (n) { UB(16) UB(32) UB(64) UB(96) UB(128) UB(192) UB(256) default: cudaFree(d); return fail(SRB_EINVAL, "srb_ubench_mma: n=%d", n); }
#undef UB
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(cudaMemcpy(out2_host, d, 16, cudaMemcpyDeviceToHost));
    cudaFree(d);
    return SRB_OK;
}

extern "C" int srb_ubench_mix(srb_ctx* ctx, int nwarps, const int* n, const int* cnt, const int* acc, int rows, int cta_pair,
                              long long* out_host) {
    if (!ctx || !n || !cnt || !acc || !out_host || nwarps < 1 || nwarps > 4) return fail(SRB_EINVAL, "srb_ubench_mix: bad argument");
    MixArgs a{};
    a.nwarps = nwarps; a.rows = rows; a.commits = 2;
    int cols = 0;
    for (int w = 0; w < nwarps; ++w) {
        a.n[w] = n[w]; a.cnt[w] = cnt[w]; a.acc[w] = acc[w] < 1 ? 1 : acc[w];
        if (n[w] % 16 || n[w] < 16 || n[w] > 256) return fail(SRB_EINVAL, "srb_ubench_mix: n=%d", n[w]);
        cols += a.acc[w] * n[w];
    }
    if (cols > 512) return fail(SRB_EINVAL, "srb_ubench_mix: %d TMEM columns", cols);
    DeviceGuard dev_guard(ctx->device);
    long long* d = nullptr;
    CUDA_TRY(cudaMalloc(&d, 16 * sizeof(long long)));
    CUDA_TRY(cudaMemset(d, 0, 16 * sizeof(long long)));
    const int smem = 130 * 1024;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(cta_pair ? 2 : 1); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = nullptr;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cta_pair ? 2 : 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (cta_pair) {
        CUDA_TRY(cudaFuncSetAttribute(ubench_mix_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CUDA_TRY(cudaLaunchKernelEx(&cfg, ubench_mix_kernel<true>, a, d));
    } else {
        CUDA_TRY(cudaFuncSetAttribute(ubench_mix_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CUDA_TRY(cudaLaunchKernelEx(&cfg, ubench_mix_kernel<false>, a, d));
    }
    CUDA_TRY(cudaDeviceSynchronize());
    CUDA_TRY(cudaMemcpy(out_host, d, 16 * sizeof(long long), cudaMemcpyDeviceToHost));
    cudaFree(d);
    return SRB_OK;
}
#endif  // SRB_DEBUG

// ------------------------------------------------------------------------------------------
// backward: weight / bias gradients (train.py:57 autograd of the convs)
// ------------------------------------------------------------------------------------------
static int encode_c8_map(srb_ctx* ctx, CUtensorMap* tm, const void* base, int N, int H, int W, int planes, int bw, int bh, int bp) {
    cuuint64_t gdim[4] = {(cuuint64_t)W * 4, (cuuint64_t)H, (cuuint64_t)planes, (cuuint64_t)N};
    cuuint64_t gstr[3] = {(cuuint64_t)W * 16, (cuuint64_t)W * H * 16, (cuuint64_t)W * H * 16 * planes};
    cuuint32_t box[4] = {(cuuint32_t)bw * 4, (cuuint32_t)bh, (cuuint32_t)bp, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = ctx->encode(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(base), gdim, gstr, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled failed: %d", (int)r);
    return SRB_OK;
}

// dY tensor map with the dimensions ordered (pixel, plane, row, image): a box lands in shared memory as [row][plane][px],
// which makes the (row shift, plane) channel groups of the folded weight-gradient GEMM equidistant (srb_wgrad.cuh)
static int encode_dy_map(srb_ctx* ctx, CUtensorMap* tm, const void* base, int N, int H, int W, int planes, int box_planes = WgradGeom::DYPLANES,
                         int box_rows = WgradGeom::DYH) {
    using G = WgradGeom;
    cuuint64_t gdim[4] = {(cuuint64_t)W * 4, (cuuint64_t)planes, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t gstr[3] = {(cuuint64_t)W * H * 16, (cuuint64_t)W * 16, (cuuint64_t)W * H * 16 * planes};
    cuuint32_t box[4] = {(cuuint32_t)G::TW * 4, (cuuint32_t)box_planes, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = ctx->encode(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(base), gdim, gstr, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled (dY, plane-minor rows) failed: %d", (int)r);
    return SRB_OK;
}

// input-channel blocks of a weight-gradient GEMM: N = block width, at most 160 (three fp32 accumulators in 512 TMEM columns)
static void wgrad_blocks(int cin, int* nblk, int* blk) {
    *nblk = (cin + 159) / 160;
    *blk = round_up((cin + *nblk - 1) / *nblk, 16);
}

// units of one conv appended to `units`; tm = index of its tensor-map pair
static int wgrad_units(const srb_grad_job& q, int tm, std::vector<WgradUnit>& units) {
    using G = WgradGeom;
    int nblk, blk;
    wgrad_blocks(q.cin, &nblk, &blk);
    WgradUnit u{};
    u.tm = tm; u.N = q.N; u.H = q.H; u.W = q.W;
    u.tiles_x = (q.W + G::TW - 1) / G::TW; u.tiles_y = (q.H + G::TH - 1) / G::TH;
    long long nt = (long long)u.tiles_x * u.tiles_y * q.N;
    if (nt <= 0 || nt > 0x7fffffffLL) return fail(SRB_EINVAL, "weight gradient: %lld tiles", nt);
    u.num_tiles = (int)nt; u.cin_total = q.cin; u.scale = q.scale; u.dw = q.dw; u.xplanes = blk / 8;
    for (int ci0 = 0; ci0 < q.cin; ci0 += blk)
        for (int co0 = 0; co0 < q.cout; co0 += 32) {
            u.cin0 = ci0; u.cout0 = co0;
            u.cin_blk = q.cin - ci0 < blk ? q.cin - ci0 : blk;
            u.cout_blk = q.cout - co0 < 32 ? q.cout - co0 : 32;
            u.x_c8_off = q.x_c8_off + ci0 / 8; u.dy_c8_off = q.dy_c8_off + co0 / 8;
            u.db = (q.db && ci0 == 0) ? q.db + co0 : nullptr;     // the bias gradient rides on the first input-channel block
            units.push_back(u);
        }
    return SRB_OK;
}

extern "C" int srb_conv2d_wgrad(srb_ctx* ctx, const void* x, int x_c8_total, int x_c8_off, int cin, const void* dy,
                                int dy_c8_total, int dy_c8_off, int cout, float* dw_oihw, float scale, int k, int dtype,
                                int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !x || !dy || !dw_oihw) return fail(SRB_EINVAL, "srb_conv2d_wgrad: NULL argument");
    if (k != 3) return fail(SRB_ENOSUP, "srb_conv2d_wgrad: tensor-core path is 3x3 (use srb_conv2d_wgrad_direct for 9x9)");
    if (cin % 8 || cout % 8) return fail(SRB_ENOSUP, "srb_conv2d_wgrad: cin=%d cout=%d must be multiples of 8", cin, cout);
    if (((uintptr_t)x & 15) || ((uintptr_t)dy & 15)) return fail(SRB_EINVAL, "srb_conv2d_wgrad: buffers must be 16-byte aligned");
    if (x_c8_off < 0 || x_c8_off + cin / 8 > x_c8_total || dy_c8_off < 0 || dy_c8_off + cout / 8 > dy_c8_total)
        return fail(SRB_EINVAL, "srb_conv2d_wgrad: plane slice out of range");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_conv2d_wgrad: empty shape");
    using G = WgradGeom;
    srb_grad_job q{};
    q.x = x; q.x_c8_total = x_c8_total; q.x_c8_off = x_c8_off; q.cin = cin; q.dy = dy; q.dy_c8_total = dy_c8_total;
    q.dy_c8_off = dy_c8_off; q.cout = cout; q.dw = dw_oihw; q.db = nullptr; q.scale = scale; q.N = N; q.H = H; q.W = W;
    std::vector<WgradUnit> units;
    int rc;
    if ((rc = wgrad_units(q, 0, units))) return rc;
    CUtensorMap tx, tdy;
    if ((rc = encode_c8_map(ctx, &tx, x, N, H, W, x_c8_total, G::XW, G::TH, units[0].xplanes))) return rc;
    if ((rc = encode_dy_map(ctx, &tdy, dy, N, H, W, dy_c8_total))) return rc;
    // one launch: units x split-K parts of their pixel tiles (red.global.add), about one CTA per SM
    int grid = ctx->num_sms / (int)units.size();
    if (grid < 1) grid = 1;
    if (grid > units[0].num_tiles) grid = units[0].num_tiles;
    WgradUnit u0 = units[0];
    u0.x_c8_off = x_c8_off; u0.dy_c8_off = dy_c8_off;
    const dim3 g2(grid, (unsigned)units.size());
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16) {
        CUDA_TRY(cudaFuncSetAttribute(wgrad_tc_kernel<true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        wgrad_tc_kernel<true, 3><<<g2, 256, G::SMEM_BYTES, st>>>(tx, tdy, u0, cin, cout, u0.xplanes * 8, ctx->dbg_skip);
    } else {
        CUDA_TRY(cudaFuncSetAttribute(wgrad_tc_kernel<false, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        wgrad_tc_kernel<false, 3><<<g2, 256, G::SMEM_BYTES, st>>>(tx, tdy, u0, cin, cout, u0.xplanes * 8, ctx->dbg_skip);
    }
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_bias_grad(srb_ctx* ctx, const void* dy, int dy_c8_total, int dy_c8_off, int cout, float* db, float scale,
                             int dtype, int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !dy || !db) return fail(SRB_EINVAL, "srb_bias_grad: NULL argument");
    int planes = (cout + 7) / 8;
    if (dy_c8_off < 0 || dy_c8_off + planes > dy_c8_total) return fail(SRB_EINVAL, "srb_bias_grad: plane slice out of range");
    size_t HW = (size_t)H * W, total = HW * N;
    int gx = (int)((total + 255) / 256); if (gx > ctx->num_sms * 2) gx = ctx->num_sms * 2; if (gx < 1) gx = 1;
    dim3 grid(gx, planes);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16) bias_grad_kernel<true><<<grid, 256, 0, st>>>((const uint4*)dy, dy_c8_total, dy_c8_off, planes, HW, N, scale, db, cout);
    else bias_grad_kernel<false><<<grid, 256, 0, st>>>((const uint4*)dy, dy_c8_total, dy_c8_off, planes, HW, N, scale, db, cout);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// backward glue + 9x9 weight gradients (srb_bwd.cuh)
// ------------------------------------------------------------------------------------------
extern "C" int srb_tanh_bwd_to_c8(srb_ctx* ctx, const float* gout, const float* out, void* g_c8, float scale, int dtype, int N,
                                  int H, int W, srb_stream_t stream) {
    if (!ctx || !gout || !out || !g_c8) return fail(SRB_EINVAL, "srb_tanh_bwd_to_c8: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_tanh_bwd_to_c8: empty shape");
    size_t HW = (size_t)H * W, total = HW * N;
    int grid = ew_grid(ctx, total, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16) tanh_bwd_to_c8_kernel<true><<<grid, 256, 0, st>>>(gout, out, (uint4*)g_c8, scale, HW, total);
    else tanh_bwd_to_c8_kernel<false><<<grid, 256, 0, st>>>(gout, out, (uint4*)g_c8, scale, HW, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_pixunshuffle_mask(srb_ctx* ctx, const void* g_u, const void* u, void* g_s, int c8_u, float slope, int dtype,
                                     int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !g_u || !u || !g_s) return fail(SRB_EINVAL, "srb_pixunshuffle_mask: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0 || c8_u <= 0) return fail(SRB_EINVAL, "srb_pixunshuffle_mask: empty shape");
    size_t total = (size_t)N * c8_u * 4 * H * W;
    int grid = ew_grid(ctx, total, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16) pixunshuffle_mask_kernel<true><<<grid, 256, 0, st>>>(g_u, u, (uint4*)g_s, c8_u, H, W, slope, total);
    else pixunshuffle_mask_kernel<false><<<grid, 256, 0, st>>>(g_u, u, (uint4*)g_s, c8_u, H, W, slope, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_pixshuffle_lrelu(srb_ctx* ctx, const void* s_c8, void* u_c8, int c8_u, float slope, int dtype, int N, int H, int W,
                                    srb_stream_t stream) {
    if (!ctx || !s_c8 || !u_c8) return fail(SRB_EINVAL, "srb_pixshuffle_lrelu: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0 || c8_u <= 0) return fail(SRB_EINVAL, "srb_pixshuffle_lrelu: empty shape");
    const size_t total = (size_t)N * c8_u * 4 * H * W;
    const int grid = ew_grid(ctx, total, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16) pixshuffle_lrelu_kernel<true><<<grid, 256, 0, st>>>(s_c8, (uint4*)u_c8, c8_u, H, W, slope, total);
    else pixshuffle_lrelu_kernel<false><<<grid, 256, 0, st>>>(s_c8, (uint4*)u_c8, c8_u, H, W, slope, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_c8_axpby_mask(srb_ctx* ctx, const void* a, int a_c8_total, int a_c8_off, float alpha, const void* b,
                                 int b_c8_total, int b_c8_off, float beta, const void* act, int act_c8_total, int act_c8_off,
                                 float slope, void* out, int out_c8_total, int out_c8_off, int planes, int dtype, int N, int H,
                                 int W, srb_stream_t stream) {
    if (!ctx || !a || !out) return fail(SRB_EINVAL, "srb_c8_axpby_mask: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0 || planes <= 0) return fail(SRB_EINVAL, "srb_c8_axpby_mask: empty shape");
    if (a_c8_off < 0 || a_c8_off + planes > a_c8_total || out_c8_off < 0 || out_c8_off + planes > out_c8_total ||
        (b && (b_c8_off < 0 || b_c8_off + planes > b_c8_total)) || (act && (act_c8_off < 0 || act_c8_off + planes > act_c8_total)))
        return fail(SRB_EINVAL, "srb_c8_axpby_mask: plane slice out of range");
    size_t HW = (size_t)H * W, total = HW * N * planes;
    int grid = ew_grid(ctx, total, 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SRB_DT_BF16)
        c8_axpby_mask_kernel<true><<<grid, 256, 0, st>>>((const uint4*)a, a_c8_total, a_c8_off, (const uint4*)b, b_c8_total, b_c8_off,
            (const uint4*)act, act_c8_total, act_c8_off, (uint4*)out, out_c8_total, out_c8_off, planes, alpha, beta, slope, HW, total);
    else
        c8_axpby_mask_kernel<false><<<grid, 256, 0, st>>>((const uint4*)a, a_c8_total, a_c8_off, (const uint4*)b, b_c8_total, b_c8_off,
            (const uint4*)act, act_c8_total, act_c8_off, (uint4*)out, out_c8_total, out_c8_off, planes, alpha, beta, slope, HW, total);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

extern "C" int srb_conv9_wgrad(srb_ctx* ctx, const void* x, int x_c8_total, int x_c8_off, int cin, const void* dy, int dy_c8_total,
                               int dy_c8_off, int cout, float* dw_oihw, float scale, int dtype, int N, int H, int W,
                               srb_stream_t stream) {
    // the folded tensor-core weight-gradient GEMM (srb_wgrad.cuh, KS = 9): units of (8 output channels = one dY plane,
    // <= 32 input channels), every unit's pixel tiles split over the SMs left per unit
    if (!ctx || !x || !dy || !dw_oihw) return fail(SRB_EINVAL, "srb_conv9_wgrad: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_conv9_wgrad: empty shape");
    if (cin <= 0 || cout <= 0) return fail(SRB_EINVAL, "srb_conv9_wgrad: cin=%d cout=%d", cin, cout);
    if (((uintptr_t)x & 15) || ((uintptr_t)dy & 15)) return fail(SRB_EINVAL, "srb_conv9_wgrad: buffers must be 16-byte aligned");
    const int xp = (cin + 7) / 8, dyp = (cout + 7) / 8;
    if (x_c8_off < 0 || x_c8_off + xp > x_c8_total || dy_c8_off < 0 || dy_c8_off + dyp > dy_c8_total)
        return fail(SRB_EINVAL, "srb_conv9_wgrad: plane slice out of range");
    using G = WgradGeomT<9>;
    const int blk = cin < 32 ? round_up(cin, 16) : 32;          // MMA N: 16 for the head's 3 input channels, else 32
    CUtensorMap tx, tdy;
    int rc;
    if ((rc = encode_c8_map(ctx, &tx, x, N, H, W, x_c8_total, G::XW, G::TH, blk / 8))) return rc;
    if ((rc = encode_dy_map(ctx, &tdy, dy, N, H, W, dy_c8_total, G::DYPLANES, G::DYH))) return rc;
    WgradUnit u{};
    u.N = N; u.H = H; u.W = W;
    u.tiles_x = (W + G::TW - 1) / G::TW; u.tiles_y = (H + G::TH - 1) / G::TH;
    long long nt = (long long)u.tiles_x * u.tiles_y * N;
    if (nt > 0x7fffffffLL) return fail(SRB_EINVAL, "srb_conv9_wgrad: %lld tiles", nt);
    u.num_tiles = (int)nt; u.cin_total = cin; u.scale = scale; u.dw = dw_oihw; u.db = nullptr; u.xplanes = blk / 8;
    const int nunits = ((cin + blk - 1) / blk) * dyp;
    int grid = ctx->num_sms / nunits;
    if (grid < 1) grid = 1;
    if (grid > u.num_tiles) grid = u.num_tiles;
    cudaStream_t st = (cudaStream_t)stream;
    u.x_c8_off = x_c8_off; u.dy_c8_off = dy_c8_off;
    const dim3 g2(grid, (unsigned)nunits);
    if (dtype == SRB_DT_BF16) {
        CUDA_TRY(cudaFuncSetAttribute(wgrad_tc_kernel<true, 9>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        wgrad_tc_kernel<true, 9><<<g2, 256, G::SMEM_BYTES, st>>>(tx, tdy, u, cin, cout, blk, ctx->dbg_skip);
    } else {
        CUDA_TRY(cudaFuncSetAttribute(wgrad_tc_kernel<false, 9>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        wgrad_tc_kernel<false, 9><<<g2, 256, G::SMEM_BYTES, st>>>(tx, tdy, u, cin, cout, blk, ctx->dbg_skip);
    }
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// batched weight / bias gradients and batched operand packing (one launch per training step each)
// ------------------------------------------------------------------------------------------
struct srb_grad_batch {
    int dtype = 0, nunits = 0, nbias = 0;
    CUtensorMap* d_tmaps = nullptr;
    WgradUnit* d_units = nullptr;
    BiasUnit* d_bias = nullptr;
};

extern "C" void srb_grad_batch_destroy(srb_grad_batch* b) {
    if (!b) return;
    if (b->d_tmaps) cudaFree(b->d_tmaps);
    if (b->d_units) cudaFree(b->d_units);
    if (b->d_bias) cudaFree(b->d_bias);
    delete b;
}

extern "C" int srb_grad_batch_create(srb_ctx* ctx, const srb_grad_job* jobs, int n, int dtype, srb_grad_batch** out) {
    if (!ctx || !jobs || !out || n <= 0) return fail(SRB_EINVAL, "srb_grad_batch_create: NULL / empty argument");
    using G = WgradGeom;
    std::vector<CUtensorMap> tmaps;
    std::vector<WgradUnit> units;
    std::vector<BiasUnit> bias;
    for (int j = 0; j < n; ++j) {
        const srb_grad_job& q = jobs[j];
        if (!q.dy || q.N <= 0 || q.H <= 0 || q.W <= 0) return fail(SRB_EINVAL, "srb_grad_batch_create: job %d: bad dY / shape", j);
        if ((uintptr_t)q.dy & 15) return fail(SRB_EINVAL, "srb_grad_batch_create: job %d: dY must be 16-byte aligned", j);
        const int dyp = (q.cout + 7) / 8;
        if (q.dy_c8_off < 0 || q.dy_c8_off + dyp > q.dy_c8_total) return fail(SRB_EINVAL, "srb_grad_batch_create: job %d: dY slice out of range", j);
        if (q.dw) {
            if (!q.x || ((uintptr_t)q.x & 15)) return fail(SRB_EINVAL, "srb_grad_batch_create: job %d: x must be a 16-byte aligned buffer", j);
            if (q.cin % 8 || q.cout % 8) return fail(SRB_ENOSUP, "srb_grad_batch_create: job %d: cin=%d cout=%d must be multiples of 8", j, q.cin, q.cout);
            if (q.x_c8_off < 0 || q.x_c8_off + q.cin / 8 > q.x_c8_total) return fail(SRB_EINVAL, "srb_grad_batch_create: job %d: x slice out of range", j);
            CUtensorMap tx, tdy;
            int rc, nblk, blk;
            wgrad_blocks(q.cin, &nblk, &blk);
            if ((rc = encode_c8_map(ctx, &tx, q.x, q.N, q.H, q.W, q.x_c8_total, G::XW, G::TH, blk / 8))) return rc;
            if ((rc = encode_dy_map(ctx, &tdy, q.dy, q.N, q.H, q.W, q.dy_c8_total))) return rc;
            const int tm = (int)tmaps.size() / 2;
            tmaps.push_back(tx); tmaps.push_back(tdy);
            if ((rc = wgrad_units(q, tm, units))) return rc;
        }
        if (q.db && !q.dw)      // bias gradients without a 3x3 weight gradient (9x9 head / tail): the reduction kernel
            for (int pl = 0; pl < dyp; ++pl) {
                BiasUnit bu{};
                bu.dy = (const uint4*)q.dy; bu.c8_total = q.dy_c8_total; bu.plane = q.dy_c8_off + pl; bu.N = q.N;
                bu.HW = (size_t)q.H * q.W; bu.scale = q.scale; bu.db = q.db + pl * 8;
                bu.nch = q.cout - pl * 8 < 8 ? q.cout - pl * 8 : 8;
                bias.push_back(bu);
            }
    }
    // longest units first: the block scheduler hands CTAs to SMs in index order as they free up
    auto unit_cost = [](const WgradUnit& u) { const int n = u.xplanes * 8; return (long long)u.num_tiles * (n / 2 > 32 + n / 4 ? n / 2 : 32 + n / 4); };
    if (!(ctx->path & SRB_PATH_UNSORTED_UNITS))
        std::stable_sort(units.begin(), units.end(), [&](const WgradUnit& a, const WgradUnit& c) { return unit_cost(a) > unit_cost(c); });
    srb_grad_batch* b = new srb_grad_batch();
    b->dtype = dtype; b->nunits = (int)units.size(); b->nbias = (int)bias.size();
    DeviceGuard dev_guard(ctx->device);
    bool ok = true;
    if (!tmaps.empty()) ok = ok && cudaMalloc(&b->d_tmaps, tmaps.size() * sizeof(CUtensorMap)) == cudaSuccess &&
        cudaMemcpy(b->d_tmaps, tmaps.data(), tmaps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice) == cudaSuccess;
    if (!units.empty()) ok = ok && cudaMalloc(&b->d_units, units.size() * sizeof(WgradUnit)) == cudaSuccess &&
        cudaMemcpy(b->d_units, units.data(), units.size() * sizeof(WgradUnit), cudaMemcpyHostToDevice) == cudaSuccess;
    if (!bias.empty()) ok = ok && cudaMalloc(&b->d_bias, bias.size() * sizeof(BiasUnit)) == cudaSuccess &&
        cudaMemcpy(b->d_bias, bias.data(), bias.size() * sizeof(BiasUnit), cudaMemcpyHostToDevice) == cudaSuccess;
    if (!ok) { srb_grad_batch_destroy(b); return fail(SRB_ENOMEM, "srb_grad_batch_create: device allocation / upload failed"); }
    *out = b;
    return SRB_OK;
}

extern "C" int srb_grad_batch_units(const srb_grad_batch* b) { return b ? b->nunits : 0; }

extern "C" int srb_grad_batch_run(srb_ctx* ctx, srb_grad_batch* b, int splitk, srb_stream_t stream) {
    if (!ctx || !b) return fail(SRB_EINVAL, "srb_grad_batch_run: NULL argument");
    if (splitk < 1) splitk = 1;
    cudaStream_t st = (cudaStream_t)stream;
    using G = WgradGeom;
    if (b->nunits) {
        if (b->dtype == SRB_DT_BF16) {
            CUDA_TRY(cudaFuncSetAttribute(wgrad_batch_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
            wgrad_batch_kernel<true><<<b->nunits * splitk, 256, G::SMEM_BYTES, st>>>(b->d_tmaps, b->d_units, splitk, ctx->dbg_skip);
        } else {
            CUDA_TRY(cudaFuncSetAttribute(wgrad_batch_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
            wgrad_batch_kernel<false><<<b->nunits * splitk, 256, G::SMEM_BYTES, st>>>(b->d_tmaps, b->d_units, splitk, ctx->dbg_skip);
        }
        CUDA_TRY(cudaGetLastError());
    }
    if (b->nbias) {
        const dim3 bg(b->nbias, 64);
        if (b->dtype == SRB_DT_BF16) bias_grad_batch_kernel<true><<<bg, 256, 0, st>>>(b->d_bias);
        else bias_grad_batch_kernel<false><<<bg, 256, 0, st>>>(b->d_bias);
        CUDA_TRY(cudaGetLastError());
    }
    return SRB_OK;
}

struct srb_pack_batch {
    int n = 0; size_t max_total = 0;
    PackJob* d_jobs = nullptr;
};

extern "C" void srb_pack_batch_destroy(srb_pack_batch* b) {
    if (!b) return;
    if (b->d_jobs) cudaFree(b->d_jobs);
    delete b;
}

extern "C" int srb_pack_batch_create(srb_ctx* ctx, const srb_pack_job* jobs, int n, srb_pack_batch** out) {
    if (!ctx || !jobs || !out || n <= 0) return fail(SRB_EINVAL, "srb_pack_batch_create: NULL / empty argument");
    std::vector<PackJob> v(n);
    size_t mx = 0;
    for (int j = 0; j < n; ++j) {
        const srb_pack_job& q = jobs[j];
        srb_weights* w = q.w;
        if (!w || !q.w_src) return fail(SRB_EINVAL, "srb_pack_batch_create: job %d: NULL", j);
        if (q.c_n < 0 || q.c_lo < 0 || (q.c_n > 0 && (q.c_lo + q.c_n > w->cin || (w->mode != MODE_STD && w->mode != MODE_DYF))))
            return fail(SRB_EINVAL, "srb_pack_batch_create: job %d: channel range [%d, +%d) of a %d-channel 3x3 operand", j, q.c_lo, q.c_n, w->cin);
        const int ncin = q.c_n > 0 ? q.c_n : w->cin;
        const int so_need = q.transpose ? q.c_off + ncin : q.o_off + w->cout;
        const int si_need = q.transpose ? q.o_off + w->cout : q.c_off + ncin;
        if (q.o_off < 0 || q.c_off < 0 || so_need > q.src_o || si_need > q.src_i)
            return fail(SRB_EINVAL, "srb_pack_batch_create: job %d: slice exceeds the source filter", j);
        PackJob& pj = v[j];
        pj.p.mode = w->mode; pj.p.c8in = w->c8in; pj.p.nout = w->nout; pj.p.nmma = w->nmma; pj.p.kblocks = w->kblocks; pj.p.cout = w->cout;
        pj.p.cin = w->cin; pj.p.k = w->k; pj.p.src_i = q.src_i; pj.p.o_off = q.o_off; pj.p.c_off = q.c_off;
        pj.p.transpose = q.transpose; pj.p.total = w->pack_bytes / 2;
        pj.p.c_lo = q.c_lo; pj.p.c_n = q.c_n; pj.p.scale = q.scale == 0.f ? 1.f : q.scale;
        pj.w = q.w_src; pj.pack = (uint16_t*)w->d_pack; pj.bias_src = q.transpose ? nullptr : q.bias_src; pj.bias_dst = w->d_bias;
        pj.bf16 = w->dtype == SRB_DT_BF16;
        if (pj.p.total > mx) mx = pj.p.total;
    }
    srb_pack_batch* b = new srb_pack_batch();
    b->n = n; b->max_total = mx;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&b->d_jobs, n * sizeof(PackJob)) != cudaSuccess ||
        cudaMemcpy(b->d_jobs, v.data(), n * sizeof(PackJob), cudaMemcpyHostToDevice) != cudaSuccess) {
        srb_pack_batch_destroy(b);
        return fail(SRB_ENOMEM, "srb_pack_batch_create: device allocation / upload failed");
    }
    *out = b;
    return SRB_OK;
}

extern "C" int srb_pack_batch_run(srb_ctx* ctx, srb_pack_batch* b, srb_stream_t stream) {
    if (!ctx || !b) return fail(SRB_EINVAL, "srb_pack_batch_run: NULL argument");
    int gx = (int)((b->max_total + 256 * 8 - 1) / (256 * 8));
    if (gx < 1) gx = 1;
    if (gx > 64) gx = 64;
    dim3 grid(gx, b->n);
    pack_conv_batch_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(b->d_jobs);
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// fused residual block (srb_resblock.cuh)
// ------------------------------------------------------------------------------------------
struct RBMaps { const void* x = nullptr; int N = 0, H = 0, W = 0; int kind = 0; CUtensorMap tm; };   // kind 1: 5-D whole-row boxes (tail)
struct srb_resblock {
    int dtype = 0;
    uint8_t* d_wA = nullptr; uint8_t* d_wB = nullptr;
    uint8_t* d_wA2 = nullptr; uint8_t* d_wB2 = nullptr;       // CTA-pair packing: [rank][4 arrangements]
    float* d_bA = nullptr; float* d_bB = nullptr;
    float h_bias[32 + 64] = {};      // the kernel takes the biases by value (constant bank)
    mutable RBMaps maps[2];          // tensor maps of the last two (x, y, shape) pairs: a trunk ping-pongs two buffers
    mutable int next_map = 0;
    mutable bool attr_set = false;
};

extern "C" void srb_resblock_destroy(srb_resblock* rb) {
    if (!rb) return;
    if (rb->d_wA) cudaFree(rb->d_wA);
    if (rb->d_wB) cudaFree(rb->d_wB);
    if (rb->d_wA2) cudaFree(rb->d_wA2);
    if (rb->d_wB2) cudaFree(rb->d_wB2);
    if (rb->d_bA) cudaFree(rb->d_bA);
    if (rb->d_bB) cudaFree(rb->d_bB);
    delete rb;
}

// fold packing: K-step ks = dx * (cin/16) + c2 ; [2 granules][3*cout rows: slice s = 2 - dy, then co][8 k]
static void pack_fold(const float* w, int cout, int cin, int dtype, std::vector<uint16_t>& out) {
    const int h = cin / 16, rows = 3 * cout;
    out.assign((size_t)3 * h * 2 * rows * 8, 0);
    for (int dx = 0; dx < 3; ++dx) for (int c2 = 0; c2 < h; ++c2) for (int g = 0; g < 2; ++g)
        for (int s = 0; s < 3; ++s) for (int co = 0; co < cout; ++co) for (int j = 0; j < 8; ++j) {
            const int ks = dx * h + c2, c = (2 * c2 + g) * 8 + j, dy = 2 - s;
            out[(((size_t)ks * 2 + g) * rows + s * cout + co) * 8 + j] = f2h(w[(((size_t)co * cin + c) * 3 + dy) * 3 + dx], dtype);
        }
}

// CTA-pair fold packing (RBGeom in srb_resblock.cuh): for rank v of the pair and each arrangement, [K-step][2 granules][rows][8]
// where the rows are this CTA's half of every piece (s0, n) of B = [W(dy=2) | W(dy=1) | W(dy=0)] (global row = s * cout + co):
// piece (s0, n) -> rows s0 * cout + v * n * cout / 2 + [0, n * cout / 2).
static void pack_fold_pair(const float* w, int cout, int cin, int dtype, std::vector<uint16_t>& out) {
    const int h = cin / 16, ksteps = 3 * h, rowsF = 3 * cout / 2, rows3 = cout / 2;
    const size_t arrF = (size_t)ksteps * 2 * rowsF * 8, arr3 = (size_t)ksteps * 2 * rows3 * 8, per_rank = 3 * arrF + arr3;
    out.assign(2 * per_rank, 0);
    struct Piece { int s0, n; };
    const Piece arrs[4][2] = {{{0, 3}, {0, 0}}, {{0, 1}, {1, 2}}, {{0, 2}, {2, 1}}, {{1, 1}, {0, 0}}};
    for (int v = 0; v < 2; ++v) for (int a = 0; a < 4; ++a) {
        const int rows = a == 3 ? rows3 : rowsF;
        uint16_t* base = out.data() + v * per_rank + (a < 3 ? a * arrF : 3 * arrF);
        std::vector<int> grow;                       // global B row of each of this CTA's rows
        for (int pc = 0; pc < 2; ++pc) {
            const Piece& q = arrs[a][pc];
            for (int i = 0; i < q.n * cout / 2; ++i) grow.push_back(q.s0 * cout + v * q.n * cout / 2 + i);
        }
        for (int dx = 0; dx < 3; ++dx) for (int c2 = 0; c2 < h; ++c2) for (int g = 0; g < 2; ++g)
            for (int r = 0; r < rows && r < (int)grow.size(); ++r) for (int j = 0; j < 8; ++j) {
                const int ks = dx * h + c2, c = (2 * c2 + g) * 8 + j, sl = grow[r] / cout, co = grow[r] % cout, dy = 2 - sl;
                base[(((size_t)ks * 2 + g) * rows + r) * 8 + j] = f2h(w[(((size_t)co * cin + c) * 3 + dy) * 3 + dx], dtype);
            }
    }
}

extern "C" int srb_resblock_create(srb_ctx* ctx, const float* wA, const float* bA, const float* wB, const float* bB, int dtype,
                                   srb_resblock** out) {
    if (!ctx || !wA || !bA || !wB || !bB || !out) return fail(SRB_EINVAL, "srb_resblock_create: NULL argument");
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "srb_resblock_create: dtype %d", dtype);
    using G = RBGeom;
    std::vector<uint16_t> pa, pb;
    pack_fold(wA, G::CH, G::CA, dtype, pa);
    pack_fold(wB, G::CB, G::CH, dtype, pb);
    if (pa.size() * 2 != (size_t)G::WA_BYTES || pb.size() * 2 != (size_t)G::WB_BYTES) return fail(SRB_EINVAL, "srb_resblock_create: pack size");
    srb_resblock* rb = new srb_resblock();
    rb->dtype = dtype;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&rb->d_wA, G::WA_BYTES) != cudaSuccess || cudaMalloc(&rb->d_wB, G::WB_BYTES) != cudaSuccess ||
        cudaMalloc(&rb->d_bA, G::CH * 4) != cudaSuccess || cudaMalloc(&rb->d_bB, G::CB * 4) != cudaSuccess) {
        srb_resblock_destroy(rb);
        return fail(SRB_ENOMEM, "srb_resblock_create: cudaMalloc failed");
    }
    CUDA_TRY(cudaMemcpy(rb->d_wA, pa.data(), G::WA_BYTES, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(rb->d_wB, pb.data(), G::WB_BYTES, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(rb->d_bA, bA, G::CH * 4, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(rb->d_bB, bB, G::CB * 4, cudaMemcpyHostToDevice));
    {
        std::vector<uint16_t> pa2, pb2;
        pack_fold_pair(wA, G::CH, G::CA, dtype, pa2);
        pack_fold_pair(wB, G::CB, G::CH, dtype, pb2);
        if (pa2.size() * 2 != (size_t)2 * G::W2_BYTES || pb2.size() * 2 != (size_t)2 * G::W2_BYTES) { srb_resblock_destroy(rb); return fail(SRB_EINVAL, "srb_resblock_create: pair pack size"); }
        if (cudaMalloc(&rb->d_wA2, 2 * G::W2_BYTES) != cudaSuccess || cudaMalloc(&rb->d_wB2, 2 * G::W2_BYTES) != cudaSuccess) {
            srb_resblock_destroy(rb);
            return fail(SRB_ENOMEM, "srb_resblock_create: cudaMalloc failed");
        }
        CUDA_TRY(cudaMemcpy(rb->d_wA2, pa2.data(), 2 * G::W2_BYTES, cudaMemcpyHostToDevice));
        CUDA_TRY(cudaMemcpy(rb->d_wB2, pb2.data(), 2 * G::W2_BYTES, cudaMemcpyHostToDevice));
    }
    memcpy(rb->h_bias, bA, G::CH * 4);
    memcpy(rb->h_bias + G::CH, bB, G::CB * 4);
    *out = rb;
    return SRB_OK;
}

// The row-fold kernels (fused block, conv1, tail) tile a row as 16 slots of 8 px: several whole image rows of narrow
// images, or overlapping 16-group strips of a wide one (rb_slot in srb_resblock.cuh) — every width is covered.
extern "C" int srb_resblock_supported(int H, int W) { return H > 0 && W > 0; }

// Tile geometry of the row-fold kernels; returns the number of strip-sets.  ov = 1: a wide image's strips overlap by one
// group on each side (the result of a strip's edge groups needs data of groups outside the tile).
template <class P>
static long long set_tile_geom(P& p, int N, int H, int W, int ov) {
    p.N = N; p.H = H; p.W = W; p.GX = (W + 7) / 8; p.OV = ov;
    if (p.GX <= 16) {
        p.TR = 16 / p.GX; p.TS = 0;
        return ((long long)N + p.TR - 1) / p.TR;
    }
    const int step = 16 - 2 * ov;
    p.TR = 0; p.TS = (p.GX - 2 * ov + step - 1) / step;
    return (long long)N * p.TS;
}

extern "C" int srb_resblock_fwd(srb_ctx* ctx, const srb_resblock* rb, const void* x, void* y, float slope, int N, int H, int W,
                                srb_stream_t stream) {
    if (!ctx || !rb || !x || !y) return fail(SRB_EINVAL, "srb_resblock_fwd: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_resblock_fwd: empty shape");
    if (x == y) return fail(SRB_EINVAL, "srb_resblock_fwd: in-place is not supported (neighbouring CTAs re-read halo rows of x)");
    if (((uintptr_t)x & 15) || ((uintptr_t)y & 15)) return fail(SRB_EINVAL, "srb_resblock_fwd: buffers must be 16-byte aligned");
    using G = RBGeom;
    ResblockParams p{};
    const long long sets = set_tile_geom(p, N, H, W, 1);
    RBMaps* m = nullptr;
    for (RBMaps& c : rb->maps)
        if (c.x == x && c.N == N && c.H == H && c.W == W) m = &c;
    if (!m) {
        m = &rb->maps[rb->next_map];
        rb->next_map ^= 1;
        m->x = nullptr;
        cuuint64_t gdim[4] = {(cuuint64_t)W * 4, (cuuint64_t)H, 8, (cuuint64_t)N};
        cuuint64_t gstr[3] = {(cuuint64_t)W * 16, (cuuint64_t)W * H * 16, (cuuint64_t)W * H * 16 * 8};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        // convA operand: one group with its halo; narrow rows (group-major tiles, RB_GM): the same group of the tile's TR images
        cuuint32_t box[4] = {40, 1, 8, (cuuint32_t)(RB_GM && p.TS == 0 ? p.TR : 1)};
        CUresult r = ctx->encode(&m->tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(x), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled failed: %d", (int)r);
        m->x = x; m->N = N; m->H = H; m->W = W;
    }
    if (sets * H > 0x7fffffffLL) return fail(SRB_EINVAL, "srb_resblock_fwd: too many rows");
    // CTA pairs (cta_group::2): two strip-sets march in lock-step on the two SMs of a TPC, every M = 256 MMA reads half of B from
    // each.  Needs at least two strip-sets; the leftover of an odd count runs against an empty (zero-filled, not stored) partner.
    const bool pair = sets >= 2 && !(ctx->path & SRB_PATH_NO_PAIR) && (ctx->num_sms % 2) == 0;
    p.total_rows = pair ? (int)((sets + 1) / 2 * H) : (int)(sets * H);
    p.wA = rb->d_wA; p.wB = rb->d_wB; p.wA2 = rb->d_wA2; p.wB2 = rb->d_wB2;
    memcpy(p.bias, rb->h_bias, sizeof(p.bias));
    p.x = x; p.y = y; p.slope = slope; p.dbg_skip = ctx->dbg_skip; p.dbg_trace = ctx->dbg_trace;
    int grid = p.total_rows / 8;
    if (grid < 1) grid = 1;
    if (pair) {                                   // grid counts CTAs: clusters of two
        if (grid > ctx->num_sms / 2) grid = ctx->num_sms / 2;
        if (ctx->max_grid > 0 && grid > (ctx->max_grid + 1) / 2) grid = (ctx->max_grid + 1) / 2;
        grid *= 2;
    } else {
        if (grid > ctx->num_sms) grid = ctx->num_sms;
        if (ctx->max_grid > 0 && grid > ctx->max_grid) grid = ctx->max_grid;
    }
    if (!rb->attr_set) {
        CUDA_TRY(cudaFuncSetAttribute(resblock_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(resblock_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(resblock_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM2_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(resblock_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM2_BYTES));
        rb->attr_set = true;
    }
    // programmatic dependent launch: this launch's prologue (barriers, TMEM, filters) overlaps the previous launch's tail
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(G::THREADS); cfg.dynamicSmemBytes = G::SMEM_BYTES;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (ctx->path & SRB_PATH_NO_PDL) ? 0 : 1;
    attr[1].id = cudaLaunchAttributeClusterDimension;
    attr[1].val.clusterDim.x = 2; attr[1].val.clusterDim.y = 1; attr[1].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = pair ? 2 : 1;
    if (pair) cfg.dynamicSmemBytes = G::SMEM2_BYTES;
    if (SRB_DBG(ctx->dbg_skip & 4096) && ctx->dbg_threads > 0) cfg.blockDim = dim3(ctx->dbg_threads);     // launch-overhead probes only
    if (SRB_DBG(ctx->dbg_skip & 4096) && ctx->dbg_smem >= 0) cfg.dynamicSmemBytes = ctx->dbg_smem;
    for (int rep = 0; rep <= SRB_DBG(ctx->dbg_repeat); ++rep) {
        if (pair) {
            if (rb->dtype == SRB_DT_BF16) CUDA_TRY(cudaLaunchKernelEx(&cfg, resblock_kernel<true, true>, m->tm, p));
            else CUDA_TRY(cudaLaunchKernelEx(&cfg, resblock_kernel<false, true>, m->tm, p));
        } else if (rb->dtype == SRB_DT_BF16) CUDA_TRY(cudaLaunchKernelEx(&cfg, resblock_kernel<true>, m->tm, p));
        else CUDA_TRY(cudaLaunchKernelEx(&cfg, resblock_kernel<false>, m->tm, p));
    }
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// row-fold 3x3 conv 64 -> 64 (conv1 of the model.pt graph)
// ------------------------------------------------------------------------------------------
struct srb_rowconv {
    int dtype = 0;
    uint8_t* d_w = nullptr;
    float h_bias[64] = {};
    mutable RBMaps maps[2];
    mutable int next_map = 0;
    mutable bool attr_set = false;
};

extern "C" void srb_rowconv_destroy(srb_rowconv* rc) {
    if (!rc) return;
    if (rc->d_w) cudaFree(rc->d_w);
    delete rc;
}

extern "C" int srb_rowconv_create(srb_ctx* ctx, const float* w, const float* bias, int dtype, srb_rowconv** out) {
    if (!ctx || !w || !bias || !out) return fail(SRB_EINVAL, "srb_rowconv_create: NULL argument");
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "srb_rowconv_create: dtype %d", dtype);
    using G = RCGeom;
    std::vector<uint16_t> pk;
    pack_fold(w, G::C, G::C, dtype, pk);
    if (pk.size() * 2 != (size_t)G::W_BYTES) return fail(SRB_EINVAL, "srb_rowconv_create: pack size");
    srb_rowconv* rc = new srb_rowconv();
    rc->dtype = dtype;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&rc->d_w, G::W_BYTES) != cudaSuccess) {
        srb_rowconv_destroy(rc);
        return fail(SRB_ENOMEM, "srb_rowconv_create: cudaMalloc failed");
    }
    CUDA_TRY(cudaMemcpy(rc->d_w, pk.data(), G::W_BYTES, cudaMemcpyHostToDevice));
    memcpy(rc->h_bias, bias, G::C * 4);
    *out = rc;
    return SRB_OK;
}

extern "C" int srb_rowconv_fwd(srb_ctx* ctx, const srb_rowconv* rc, const void* x, const void* res, void* y, float slope, int flags,
                               int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !rc || !x || !y) return fail(SRB_EINVAL, "srb_rowconv_fwd: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_rowconv_fwd: empty shape");
    if (x == y) return fail(SRB_EINVAL, "srb_rowconv_fwd: in-place is not supported (neighbouring CTAs re-read halo rows of x)");
    if (((uintptr_t)x & 15) || ((uintptr_t)y & 15) || ((uintptr_t)res & 15)) return fail(SRB_EINVAL, "srb_rowconv_fwd: buffers must be 16-byte aligned");
    if (flags & ~SRB_F_LRELU) return fail(SRB_ENOSUP, "srb_rowconv_fwd: flags 0x%x (only SRB_F_LRELU)", flags);
    using G = RCGeom;
    RowconvParams p{};
    const long long sets = set_tile_geom(p, N, H, W, 0);
    RBMaps* m = nullptr;
    for (RBMaps& c : rc->maps)
        if (c.x == x && c.N == N && c.H == H && c.W == W) m = &c;
    if (!m) {
        m = &rc->maps[rc->next_map];
        rc->next_map ^= 1;
        m->x = nullptr;
        cuuint64_t gdim[4] = {(cuuint64_t)W * 4, (cuuint64_t)H, 8, (cuuint64_t)N};
        cuuint64_t gstr[3] = {(cuuint64_t)W * 16, (cuuint64_t)W * H * 16, (cuuint64_t)W * H * 16 * 8};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        // narrow rows (group-major tiles): the same group of the tile's TR images in one box
        m->kind = (p.TS == 0 && !(ctx->path & SRB_PATH_GROUP_BOXES)) ? 1 : 0;
        cuuint32_t box[4] = {40, 1, 8, (cuuint32_t)(m->kind ? p.TR : 1)};
        CUresult r = ctx->encode(&m->tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(x), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled failed: %d", (int)r);
        m->x = x; m->N = N; m->H = H; m->W = W;
    }
    p.gm = m->kind;
    if (sets * H > 0x7fffffffLL) return fail(SRB_EINVAL, "srb_rowconv_fwd: too many rows");
    p.total_rows = (int)(sets * H);
    p.w = rc->d_w;
    memcpy(p.bias, rc->h_bias, sizeof(p.bias));
    p.res = res; p.y = y; p.slope = slope; p.lrelu = (flags & SRB_F_LRELU) ? 1 : 0; p.dbg_skip = ctx->dbg_skip;
    int grid = p.total_rows / 8;
    if (grid < 1) grid = 1;
    if (grid > ctx->num_sms) grid = ctx->num_sms;
    if (ctx->max_grid > 0 && grid > ctx->max_grid) grid = ctx->max_grid;
    if (!rc->attr_set) {
        CUDA_TRY(cudaFuncSetAttribute(rowconv_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(rowconv_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        rc->attr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(G::THREADS); cfg.dynamicSmemBytes = G::SMEM_BYTES;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (ctx->path & SRB_PATH_NO_PDL) ? 0 : 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (rc->dtype == SRB_DT_BF16) CUDA_TRY(cudaLaunchKernelEx(&cfg, rowconv_kernel<true>, m->tm, p));
    else CUDA_TRY(cudaLaunchKernelEx(&cfg, rowconv_kernel<false>, m->tm, p));
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// row-fold 9x9 conv 64 -> 3 + tanh -> uint8 (conv2 + TanhToArrayImage of the model.pt graph)
// ------------------------------------------------------------------------------------------
struct srb_tail9 {
    int dtype = 0;
    uint8_t* d_w = nullptr;
    float h_bias[3] = {};
    mutable RBMaps maps[2];
    mutable int next_map = 0;
    mutable bool attr_set = false;
    mutable RBMaps fmaps[2];         // srb_c1tail_fwd: tensor maps of conv1's input
    mutable int next_fmap = 0;
    mutable bool fattr_set = false;
};

extern "C" void srb_tail9_destroy(srb_tail9* t) {
    if (!t) return;
    if (t->d_w) cudaFree(t->d_w);
    delete t;
}

extern "C" int srb_tail9_create(srb_ctx* ctx, const float* w, const float* bias, int dtype, srb_tail9** out) {
    if (!ctx || !w || !bias || !out) return fail(SRB_EINVAL, "srb_tail9_create: NULL argument");
    if (dtype != SRB_DT_F16 && dtype != SRB_DT_BF16) return fail(SRB_EINVAL, "srb_tail9_create: dtype %d", dtype);
    using G = TLGeom;
    // K-step kk = 16 input channels; [2 granules][288 rows: slice s = 8 - dy, then dx * 3 + co (27 of 32)][8 k]
    std::vector<uint16_t> pk((size_t)G::W_BYTES / 2, 0);
    for (int kk = 0; kk < G::KS; ++kk) for (int g = 0; g < 2; ++g) for (int s = 0; s < 9; ++s) for (int dx = 0; dx < 9; ++dx)
        for (int co = 0; co < 3; ++co) for (int j = 0; j < 8; ++j) {
            const int c = (2 * kk + g) * 8 + j, dy = 8 - s;
            pk[(((size_t)kk * 2 + g) * G::BROWS + s * G::NS + dx * 3 + co) * 8 + j] = f2h(w[(((size_t)co * 64 + c) * 9 + dy) * 9 + dx], dtype);
        }
    srb_tail9* t = new srb_tail9();
    t->dtype = dtype;
    DeviceGuard dev_guard(ctx->device);
    if (cudaMalloc(&t->d_w, G::W_BYTES) != cudaSuccess) {
        srb_tail9_destroy(t);
        return fail(SRB_ENOMEM, "srb_tail9_create: cudaMalloc failed");
    }
    CUDA_TRY(cudaMemcpy(t->d_w, pk.data(), G::W_BYTES, cudaMemcpyHostToDevice));
    memcpy(t->h_bias, bias, 3 * 4);
    *out = t;
    return SRB_OK;
}

extern "C" int srb_tail9_fwd(srb_ctx* ctx, const srb_tail9* t, const void* x, void* y, int flags, int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !t || !x || !y) return fail(SRB_EINVAL, "srb_tail9_fwd: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_tail9_fwd: empty shape");
    if ((uintptr_t)x & 15) return fail(SRB_EINVAL, "srb_tail9_fwd: x must be 16-byte aligned");
    const int fmt = flags & (SRB_F_U8OUT | SRB_F_F32OUT);
    if ((flags & ~(SRB_F_TANH | SRB_F_U8OUT | SRB_F_F32OUT | SRB_F_HWC_BGR)) || (fmt != SRB_F_U8OUT && fmt != SRB_F_F32OUT) ||
        ((flags & SRB_F_HWC_BGR) && fmt != SRB_F_U8OUT))
        return fail(SRB_ENOSUP, "srb_tail9_fwd: flags 0x%x (SRB_F_TANH, exactly one of SRB_F_U8OUT / SRB_F_F32OUT, SRB_F_HWC_BGR only with U8OUT)", flags);
    using G = TLGeom;
    TailParams p{};
    const long long sets = set_tile_geom(p, N, H, W, 1);
    RBMaps* m = nullptr;
    for (RBMaps& c : t->maps)
        if (c.x == x && c.N == N && c.H == H && c.W == W) m = &c;
    if (!m) {
        m = &t->maps[t->next_map];
        t->next_map ^= 1;
        m->x = nullptr;
        m->kind = 0;
        CUresult r = CUDA_ERROR_INVALID_VALUE;
        if ((W % 8) == 0 && !(ctx->path & SRB_PATH_GROUP_BOXES)) {
            // whole-row boxes: the groups of a row as a tensor dimension of their own (stride 128 B), so that ONE box
            // [8 px][8 planes][groups][1 row][images] fills all 16 slots of a stage in the slot order of rb_slot
            cuuint64_t gdim[5] = {32, 8, (cuuint64_t)(W / 8), (cuuint64_t)H, (cuuint64_t)N};
            cuuint64_t gstr[4] = {(cuuint64_t)W * H * 16, 128, (cuuint64_t)W * 16, (cuuint64_t)W * H * 16 * 8};
            cuuint32_t estr[5] = {1, 1, 1, 1, 1};
            cuuint32_t box[5] = {32, 8, (cuuint32_t)(p.TS == 0 ? p.GX : 16), 1, (cuuint32_t)(p.TS == 0 ? p.TR : 1)};
            r = ctx->encode(&m->tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 5, const_cast<void*>(x), gdim, gstr, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r == CUDA_SUCCESS) m->kind = 1;
        }
        if (r != CUDA_SUCCESS) {
            cuuint64_t gdim[4] = {(cuuint64_t)W * 4, (cuuint64_t)H, 8, (cuuint64_t)N};
            cuuint64_t gstr[3] = {(cuuint64_t)W * 16, (cuuint64_t)W * H * 16, (cuuint64_t)W * H * 16 * 8};
            cuuint32_t estr[4] = {1, 1, 1, 1};
            cuuint32_t box[4] = {32, 1, 8, 1};                              // one group of 8 px, no halo
            r = ctx->encode(&m->tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(x), gdim, gstr, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        }
        if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled failed: %d", (int)r);
        m->x = x; m->N = N; m->H = H; m->W = W;
    }
    p.row5d = m->kind;
    if (sets * H > 0x7fffffffLL) return fail(SRB_EINVAL, "srb_tail9_fwd: too many rows");
    p.total_rows = (int)(sets * H);
    p.w = t->d_w;
    memcpy(p.bias, t->h_bias, sizeof(p.bias));
    p.y = y; p.flags = flags; p.dbg_skip = ctx->dbg_skip;
    int grid = p.total_rows / 8;
    if (grid < 1) grid = 1;
    if (grid > ctx->num_sms) grid = ctx->num_sms;
    if (ctx->max_grid > 0 && grid > ctx->max_grid) grid = ctx->max_grid;
    if (!t->attr_set) {
        CUDA_TRY(cudaFuncSetAttribute(tail9_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(tail9_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        t->attr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(G::THREADS); cfg.dynamicSmemBytes = G::SMEM_BYTES;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (ctx->path & SRB_PATH_NO_PDL) ? 0 : 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (t->dtype == SRB_DT_BF16) CUDA_TRY(cudaLaunchKernelEx(&cfg, tail9_kernel<true>, m->tm, p));
    else CUDA_TRY(cudaLaunchKernelEx(&cfg, tail9_kernel<false>, m->tm, p));
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// conv1 + long skip + conv2 9x9 + tanh -> uint8 in one launch (srb_c1tail.cuh): conv1's result stays on chip
// ------------------------------------------------------------------------------------------
extern "C" int srb_c1tail_supported(int H, int W) { return H > 0 && W >= 57; }     // (the exchange buffer holds rows of >= 8 groups)

extern "C" int srb_c1tail_fwd(srb_ctx* ctx, const srb_rowconv* rc, const srb_tail9* t, const void* x, const void* res, void* y, int flags,
                              int N, int H, int W, srb_stream_t stream) {
    if (!ctx || !rc || !t || !x || !res || !y) return fail(SRB_EINVAL, "srb_c1tail_fwd: NULL argument");
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_c1tail_fwd: empty shape");
    if (!srb_c1tail_supported(H, W)) return fail(SRB_ENOSUP, "srb_c1tail_fwd: W=%d (rows of at least 57 pixels; use srb_rowconv_fwd + srb_tail9_fwd)", W);
    if (rc->dtype != t->dtype) return fail(SRB_EINVAL, "srb_c1tail_fwd: conv1 and conv2 were packed for different operand types");
    if (((uintptr_t)x & 15) || ((uintptr_t)res & 15)) return fail(SRB_EINVAL, "srb_c1tail_fwd: buffers must be 16-byte aligned");
    const int fmt = flags & (SRB_F_U8OUT | SRB_F_F32OUT);
    if ((flags & ~(SRB_F_TANH | SRB_F_U8OUT | SRB_F_F32OUT | SRB_F_HWC_BGR)) || (fmt != SRB_F_U8OUT && fmt != SRB_F_F32OUT) ||
        ((flags & SRB_F_HWC_BGR) && fmt != SRB_F_U8OUT))
        return fail(SRB_ENOSUP, "srb_c1tail_fwd: flags 0x%x (SRB_F_TANH, exactly one of SRB_F_U8OUT / SRB_F_F32OUT, SRB_F_HWC_BGR only with U8OUT)", flags);
    using G = CTGeom;
    C1TailParams p{};
    const long long sets = set_tile_geom(p, N, H, W, 1);
    RBMaps* m = nullptr;
    for (RBMaps& c : t->fmaps)
        if (c.x == x && c.N == N && c.H == H && c.W == W) m = &c;
    if (!m) {
        m = &t->fmaps[t->next_fmap];
        t->next_fmap ^= 1;
        m->x = nullptr;
        cuuint64_t gdim[4] = {(cuuint64_t)W * 4, (cuuint64_t)H, 8, (cuuint64_t)N};
        cuuint64_t gstr[3] = {(cuuint64_t)W * 16, (cuuint64_t)W * H * 16, (cuuint64_t)W * H * 16 * 8};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        cuuint32_t box[4] = {40, 1, 8, 1};                              // one group with its halo
        CUresult r = ctx->encode(&m->tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(x), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "cuTensorMapEncodeTiled failed: %d", (int)r);
        m->x = x; m->N = N; m->H = H; m->W = W;
    }
    if (sets * H > 0x7fffffffLL) return fail(SRB_EINVAL, "srb_c1tail_fwd: too many rows");
    p.total_rows = (int)(sets * H);
    p.w1 = rc->d_w; p.w2 = t->d_w;
    memcpy(p.bias1, rc->h_bias, sizeof(p.bias1));
    memcpy(p.bias2, t->h_bias, sizeof(p.bias2));
    p.res = res; p.y = y; p.flags = flags;
    int grid = p.total_rows / 8;
    if (grid < 1) grid = 1;
    if (grid > ctx->num_sms) grid = ctx->num_sms;
    if (ctx->max_grid > 0 && grid > ctx->max_grid) grid = ctx->max_grid;
    if (!t->fattr_set) {
        CUDA_TRY(cudaFuncSetAttribute(c1tail_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        CUDA_TRY(cudaFuncSetAttribute(c1tail_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES));
        t->fattr_set = true;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(G::THREADS); cfg.dynamicSmemBytes = G::SMEM_BYTES;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (ctx->path & SRB_PATH_NO_PDL) ? 0 : 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (t->dtype == SRB_DT_BF16) CUDA_TRY(cudaLaunchKernelEx(&cfg, c1tail_kernel<true>, m->tm, p));
    else CUDA_TRY(cudaLaunchKernelEx(&cfg, c1tail_kernel<false>, m->tm, p));
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// Allocation-free L1 loss (srb_l1_loss_fwd_noalloc).  No counterpart in the reference: added
// for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// srb_l1_loss_fwd takes its reduction scratch (`partials`) from the caller; this variant keeps it in a file-scope device
// array, so the caller passes only inputs and outputs and the call performs no cudaMalloc / cudaFree / allocating routine.
//   * every GPU that loads the library holds the array (16.6 KB) until exit;
//   * one partial sum per 16 384-element chunk (one CTA each, summed in index order by the last CTA to finish:
//     deterministic) => at most SRB_L1_NOALLOC_MAX_ELEMS = 4096 x 16384 elements per call, larger inputs are refused;
//   * ALL calls on a device share the array, which keeps the last call's contents: the arrival counter is cleared on the
//     stream before every launch and the partials are fully rewritten, but two calls that overlap on the GPU (different
//     streams) race — serialise them.
// In host code `g_srb_noalloc_scratch` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, taken once per device with that device current.
// ------------------------------------------------------------------------------------------
constexpr int L1NA_MAX_PARTIALS = 4096;
constexpr int L1NA_CHUNK = 16384;                       // elements per CTA: 256 threads x 16 float4
static_assert((size_t)L1NA_MAX_PARTIALS * L1NA_CHUNK == SRB_L1_NOALLOC_MAX_ELEMS, "header constant");
static __device__ __align__(256) float g_srb_noalloc_scratch[L1NA_MAX_PARTIALS + 64];   // [0, 4096) partials, [4096] arrival counter

template <int DT>
__global__ void __launch_bounds__(256) l1_fwd_noalloc_kernel(const void* __restrict__ pred, const float* __restrict__ tgt, size_t n,
                                                             float scale /* w / n */, float* __restrict__ loss, float* __restrict__ scratch) {
    const size_t lo = (size_t)blockIdx.x * L1NA_CHUNK;
    const size_t hi = lo + L1NA_CHUNK < n ? lo + L1NA_CHUNK : n;
    float acc = 0.f;
    const bool vec = ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(tgt)) & 15) == 0;
    size_t i0 = lo;
    if (vec) {                                          // whole float4 groups of the chunk (lo is a multiple of 4)
        const size_t g_end = lo / 4 + (hi - lo) / 4;
        for (size_t g = lo / 4 + threadIdx.x; g < g_end; g += 256) {
            const float4 t = reinterpret_cast<const float4*>(tgt)[g];
            acc += fabsf(ldp<DT>(pred, 4 * g) - t.x) + fabsf(ldp<DT>(pred, 4 * g + 1) - t.y) + fabsf(ldp<DT>(pred, 4 * g + 2) - t.z) +
                   fabsf(ldp<DT>(pred, 4 * g + 3) - t.w);
        }
        i0 = g_end * 4;
    }
    for (size_t i = i0 + threadIdx.x; i < hi; i += 256) acc += fabsf(ldp<DT>(pred, i) - tgt[i]);
    __shared__ float s_part[8];
    __shared__ bool s_last;
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    unsigned int* counter = reinterpret_cast<unsigned int*>(scratch + L1NA_MAX_PARTIALS);
    if (threadIdx.x < 32) {
        float v = threadIdx.x < 8 ? s_part[threadIdx.x] : 0.f;
        v = warp_sum(v);
        if (threadIdx.x == 0) {
            scratch[blockIdx.x] = v;
            __threadfence();
            s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
        }
    }
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        __threadfence();
        double tot = 0.0;
        for (unsigned int b = threadIdx.x; b < gridDim.x; b += 32) tot += (double)reinterpret_cast<volatile float*>(scratch)[b];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
        if (threadIdx.x == 0) { *loss = (float)(tot * (double)scale); *counter = 0u; }
    }
}

static float* s_noalloc_scratch_dev[64] = {};           // device address of g_srb_noalloc_scratch, per device

extern "C" int srb_l1_loss_fwd_noalloc(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                                       float* loss_out, srb_stream_t stream) {
    if (!ctx || !pred || !tgt || !loss_out) return fail(SRB_EINVAL, "srb_l1_loss_fwd_noalloc: NULL argument");
    if (n == 0) return fail(SRB_EINVAL, "srb_l1_loss_fwd_noalloc: empty input (mean of nothing)");
    if (n > (size_t)SRB_L1_NOALLOC_MAX_ELEMS)
        return fail(SRB_EINVAL, "srb_l1_loss_fwd_noalloc: %zu elements exceed the built-in scratch (at most %zu = %d partial sums x %d "
                    "elements); use srb_l1_loss_fwd with caller-provided partials", n, (size_t)SRB_L1_NOALLOC_MAX_ELEMS, L1NA_MAX_PARTIALS, L1NA_CHUNK);
    if (pred_dtype < 0 || pred_dtype > 2) return fail(SRB_EINVAL, "srb_l1_loss_fwd_noalloc: pred_dtype %d", pred_dtype);
    if (ctx->device < 0 || ctx->device >= 64) return fail(SRB_ENOSUP, "srb_l1_loss_fwd_noalloc: device index %d", ctx->device);
    DeviceGuard dev_guard(ctx->device);
    float* scratch = s_noalloc_scratch_dev[ctx->device];
    if (!scratch) {                                     // set-up, once per device, with that device current
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_srb_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(SRB_ECUDA, "srb_l1_loss_fwd_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        scratch = s_noalloc_scratch_dev[ctx->device] = static_cast<float*>(p);
    }
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps the previous call's contents: the arrival counter must start at zero (an aborted launch could leave it set)
    CUDA_TRY(cudaMemsetAsync(scratch + L1NA_MAX_PARTIALS, 0, 64 * sizeof(float), st));
    const int grid = (int)((n + L1NA_CHUNK - 1) / L1NA_CHUNK);
    const float scale = (float)((double)w / (double)n);
    switch (pred_dtype) {
        case 0: l1_fwd_noalloc_kernel<0><<<grid, 256, 0, st>>>(pred, tgt, n, scale, loss_out, scratch); break;
        case 1: l1_fwd_noalloc_kernel<1><<<grid, 256, 0, st>>>(pred, tgt, n, scale, loss_out, scratch); break;
        default: l1_fwd_noalloc_kernel<2><<<grid, 256, 0, st>>>(pred, tgt, n, scale, loss_out, scratch); break;
    }
    CUDA_TRY(cudaGetLastError());
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// Rectangle copies between a host image and a device tile (rs.py:92 `window_img.to(device)`, :107 `frame.cpu()`): `planes`
// 2-D copies of width x height bytes, one cudaMemcpy2DAsync per plane, enqueued on the caller's stream.  With pinned host
// memory the tile (+ halo) goes straight from the image to the device and the interior of the result straight into the stitched
// host image — no host-side crop / paste copies and no whole-image transfers (single-image sharding).
// ------------------------------------------------------------------------------------------
extern "C" int srb_copy_rect_u8(srb_ctx* ctx, void* dst, size_t dst_pitch, size_t dst_plane, const void* src, size_t src_pitch,
                                size_t src_plane, int planes, size_t width, size_t height, int kind, srb_stream_t stream) {
    if (!ctx || !dst || !src) return fail(SRB_EINVAL, "srb_copy_rect_u8: NULL argument");
    if (planes <= 0 || width == 0 || height == 0) return fail(SRB_EINVAL, "srb_copy_rect_u8: empty rectangle");
    if (width > dst_pitch || width > src_pitch) return fail(SRB_EINVAL, "srb_copy_rect_u8: width %zu exceeds a pitch (%zu, %zu)", width, dst_pitch, src_pitch);
    cudaMemcpyKind k;
    if (kind == SRB_COPY_H2D) k = cudaMemcpyHostToDevice;
    else if (kind == SRB_COPY_D2H) k = cudaMemcpyDeviceToHost;
    else if (kind == SRB_COPY_D2D) k = cudaMemcpyDeviceToDevice;
    else return fail(SRB_EINVAL, "srb_copy_rect_u8: kind %d", kind);
    DeviceGuard dev_guard(ctx->device);
    for (int p = 0; p < planes; ++p)
        CUDA_TRY(cudaMemcpy2DAsync(static_cast<uint8_t*>(dst) + (size_t)p * dst_plane, dst_pitch,
                                   static_cast<const uint8_t*>(src) + (size_t)p * src_plane, src_pitch, width, height, k,
                                   (cudaStream_t)stream));
    return SRB_OK;
}

// ------------------------------------------------------------------------------------------
// conv chains: the weight-streaming 3x3 convs (Cout = 32) of one dense block — RDB.forward conv1 .. conv3, the gathered
// data-gradient stages of its backward — as ONE launch of persistent CTAs with tile-level dependencies (srb_conv.cuh: CHAIN)
// ------------------------------------------------------------------------------------------
struct srb_conv_chain {
    int n = 0, dtype = 0, grid = 0, num_tiles = 0;
    bool small = false, bwd = false, fold = false, wide = false;      // wide: a 64-output layer (NOUT = 64 kernel)
    ConvParams* d_layers = nullptr;
    CUtensorMap* d_tmaps = nullptr;
    unsigned int* d_flags = nullptr;      // [n][num_tiles] + epoch + ticket
};

extern "C" void srb_conv_chain_destroy(srb_conv_chain* c) {
    if (!c) return;
    if (c->d_layers) cudaFree(c->d_layers);
    if (c->d_tmaps) cudaFree(c->d_tmaps);
    if (c->d_flags) cudaFree(c->d_flags);
    delete c;
}

extern "C" int srb_conv_chain_create(srb_ctx* ctx, int n, const srb_weights* const* ws, const srb_conv_args* args, srb_conv_chain** out) {
    if (!ctx || !ws || !args || !out) return fail(SRB_EINVAL, "srb_conv_chain_create: NULL argument");
    if (n < 1 || n > CHAIN_MAX_LAYERS) return fail(SRB_EINVAL, "srb_conv_chain_create: %d layers (1 .. %d)", n, CHAIN_MAX_LAYERS);
    const int N = args[0].N, H = args[0].H, W = args[0].W;
    if (N <= 0 || H <= 0 || W <= 0) return fail(SRB_EINVAL, "srb_conv_chain_create: empty shape");
    const long long mtiles = (long long)((W + 7) / 8) * ((H + 15) / 16) * N;
    const bool fold = ws[0] && ws[0]->mode == MODE_DYF;
    const bool small = fold || (mtiles < (long long)ctx->num_sms * 6 && !(ctx->path & SRB_PATH_LARGE_TILING));   // (dy-folded: NT = 2 only)
    const int NT = small ? 2 : 4, SUPER_W = 8 * NT, SUPER_H = fold ? 14 : 16, HALO_W = SUPER_W + 2, HALO_H = SUPER_H + 2;
    std::vector<ConvParams> lay(n);
    std::vector<CUtensorMap> tms(n);
    bool bwd = false, wide = false;
    for (int l = 0; l < n; ++l) {
        const srb_weights* w = ws[l];
        const srb_conv_args& a = args[l];
        if (!w || !a.x || !a.y) return fail(SRB_EINVAL, "srb_conv_chain_create: layer %d: NULL", l);
        const bool n_ok = w->nout == 32 || (w->nout == 64 && w->mode == MODE_STD);
        if ((w->mode != MODE_STD && w->mode != MODE_DYF) || w->mode != ws[0]->mode || !w->kblocks || !n_ok || w->dtype != ws[0]->dtype)
            return fail(SRB_ENOSUP, "srb_conv_chain_create: layer %d: chains hold weight-streaming 3x3 convs with 32 (nine-tap kernels: or 64) outputs (cin=%d cout=%d)", l, w->cin, w->cout);
        wide = wide || w->nout == 64;
        if (a.N != N || a.H != H || a.W != W) return fail(SRB_EINVAL, "srb_conv_chain_create: layer %d: all layers share one shape", l);
        if (((uintptr_t)a.x & 15) || ((uintptr_t)a.y & 15) || ((uintptr_t)a.mask & 15)) return fail(SRB_EINVAL, "srb_conv_chain_create: layer %d: alignment", l);
        if (a.flags & ~(SRB_F_LRELU | SRB_F_MASK | SRB_F_RESIDUAL)) return fail(SRB_ENOSUP, "srb_conv_chain_create: layer %d: flags 0x%x (LRELU / MASK / RESIDUAL only)", l, a.flags);
        const int oplanes = (w->cout + 7) / 8;
        if (a.x_c8_off < 0 || a.x_c8_off + w->c8in > a.x_c8_total || a.y_c8_off < 0 || a.y_c8_off + oplanes > a.y_c8_total)
            return fail(SRB_EINVAL, "srb_conv_chain_create: layer %d: plane slice out of range", l);
        // residual terms (the block's last conv: add_rate * conv + x, the RRDB skip; its backward: + s_res * gy): plain loads in the epilogue
        const bool has_res = a.residual != nullptr || (a.flags & SRB_F_RESIDUAL);
        if (has_res && (!a.residual || ((uintptr_t)a.residual & 15) || a.res_c8_off < 0 ||
                        a.res_c8_off + (a.res_nplanes > 0 ? a.res_nplanes : oplanes) > a.res_c8_total))
            return fail(SRB_EINVAL, "srb_conv_chain_create: layer %d: residual", l);
        if (a.res2 && (!has_res || ((uintptr_t)a.res2 & 15) || a.res2_c8_off < 0 || a.res2_c8_off + oplanes > a.res2_c8_total))
            return fail(SRB_EINVAL, "srb_conv_chain_create: layer %d: second residual", l);
        if ((a.flags & SRB_F_MASK) && (!a.mask || a.mask_p0 < 0 || a.mask_p1 < a.mask_p0 || a.mask_c8_off + a.mask_p1 > a.mask_c8_total))
            return fail(SRB_EINVAL, "srb_conv_chain_create: layer %d: mask", l);
        bwd = bwd || (a.flags & SRB_F_MASK);
        ConvParams& p = lay[l];
        memset(&p, 0, sizeof(p));
        p.N = N; p.H = H; p.W = W;
        p.tiles_x = (W + SUPER_W - 1) / SUPER_W; p.tiles_y = (H + SUPER_H - 1) / SUPER_H;
        const long long nt = (long long)p.tiles_x * p.tiles_y * N;
        if (nt > 0x7fffffffLL) return fail(SRB_EINVAL, "srb_conv_chain_create: %lld tiles", nt);
        p.num_tiles = (int)nt;
        p.x_c8_off = a.x_c8_off; p.wpack = w->d_pack; p.bias = w->d_bias;
        p.y = a.y; p.y_c8_total = a.y_c8_total; p.y_c8_off = a.y_c8_off;
        p.res_scale = a.res_scale; p.out_scale = a.out_scale; p.slope = a.lrelu_slope;
        p.res_nplanes = a.res_nplanes > 0 ? a.res_nplanes : 1 << 30;
        if (has_res) { p.res = a.residual; p.res_c8_total = a.res_c8_total; p.res_c8_off = a.res_c8_off; }
        if (a.res2) { p.res2 = a.res2; p.res2_c8_total = a.res2_c8_total; p.res2_c8_off = a.res2_c8_off; p.res2_scale = a.res2_scale; }
        p.nmma = w->nmma;
        p.mask = a.mask; p.mask_c8_total = a.mask_c8_total; p.mask_c8_off = a.mask_c8_off;
        p.mask_p0 = a.mask_p0; p.mask_p1 = a.mask_p1; p.mask_slope = a.mask_slope;
        p.kblocks = w->kblocks; p.flags = a.flags | (has_res ? SRB_F_RESIDUAL : 0); p.cout = w->cout;
        cuuint64_t gdim[4] = {(cuuint64_t)W * 4, (cuuint64_t)H, (cuuint64_t)a.x_c8_total, (cuuint64_t)N};
        cuuint64_t gstr[3] = {(cuuint64_t)W * 16, (cuuint64_t)W * H * 16, (cuuint64_t)W * H * 16 * a.x_c8_total};
        cuuint32_t box[4] = {(cuuint32_t)HALO_W * 4, (cuuint32_t)HALO_H, 4, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        CUresult r = ctx->encode(&tms[l], CU_TENSOR_MAP_DATA_TYPE_UINT32, 4, const_cast<void*>(a.x), gdim, gstr, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(SRB_ECUDA, "srb_conv_chain_create: cuTensorMapEncodeTiled failed: %d", (int)r);
    }
    srb_conv_chain* c = new srb_conv_chain();
    c->n = n; c->dtype = ws[0]->dtype; c->num_tiles = lay[0].num_tiles; c->small = small; c->bwd = bwd; c->fold = fold; c->wide = wide;
    c->grid = c->num_tiles < ctx->num_sms ? c->num_tiles : ctx->num_sms;       // every CTA resident: the tile-level waits rely on it
    DeviceGuard dev_guard(ctx->device);
    const size_t fl = ((size_t)n * c->num_tiles + 2) * sizeof(unsigned int);
    if (cudaMalloc(&c->d_layers, n * sizeof(ConvParams)) != cudaSuccess || cudaMalloc(&c->d_tmaps, n * sizeof(CUtensorMap)) != cudaSuccess ||
        cudaMalloc(&c->d_flags, fl) != cudaSuccess) { srb_conv_chain_destroy(c); return fail(SRB_ENOMEM, "srb_conv_chain_create: cudaMalloc failed"); }
    CUDA_TRY(cudaMemcpy(c->d_layers, lay.data(), n * sizeof(ConvParams), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(c->d_tmaps, tms.data(), n * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemset(c->d_flags, 0, fl));
    *out = c;
    return SRB_OK;
}

template <int MODE, int NT, bool BF16, bool BWD, int NOUT = 32>
static int launch_chain(srb_ctx* ctx, const srb_conv_chain* c, cudaStream_t st) {
    // (KACC = 2 — two accumulators per M-tile, K-steps alternating — is implemented in conv_body and measured equal in time; it changes
    // the fp32 summation order, and chains are kept bit-identical to the same layers launched one by one)
    constexpr int KACC = 1;
    using G = ConvGeom<4, NOUT, 3, 3, MODE, NT, false, true, KACC>;
    ChainCtl cc{};
    cc.layers = c->d_layers; cc.tmaps = c->d_tmaps; cc.nlayers = c->n;
    cc.flags = c->d_flags; cc.epoch = c->d_flags + (size_t)c->n * c->num_tiles; cc.ticket = cc.epoch + 1;
    auto kern = conv_chain_kernel<4, NOUT, MODE, NT, BF16, BWD, KACC>;
    const int smem = G::SMEM_BYTES + (MODE == MODE_DYF ? 0 : CHAIN_SMEM_EXTRA);      // (the dy-folded geometry already holds the chain area)
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(c->grid); cfg.blockDim = dim3(CONV_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = (ctx->pdl || ctx->pdl_all) ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, cc));
    return SRB_OK;
}

extern "C" int srb_conv_chain_run(srb_ctx* ctx, srb_conv_chain* c, srb_stream_t stream) {
    if (!ctx || !c) return fail(SRB_EINVAL, "srb_conv_chain_run: NULL argument");
    if (c->grid > ctx->num_sms) return fail(SRB_EINVAL, "srb_conv_chain_run: grid %d exceeds the device's %d SMs", c->grid, ctx->num_sms);
    cudaStream_t st = (cudaStream_t)stream;
    const bool bf = c->dtype == SRB_DT_BF16;
#ifdef SRB_DEBUG
    for (int l = 0; l < c->n; ++l)      // probe build: the clock trace of CTA 0 (srb_debug_trace) follows the chain through its layers
        CUDA_TRY(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(c->d_layers + l) + offsetof(ConvParams, dbg_trace), &ctx->dbg_trace, sizeof(long long*),
                                 cudaMemcpyHostToDevice, st));
#endif
#define SRB_CH(MD, NTV) { if (c->bwd) return bf ? launch_chain<MD, NTV, true, true>(ctx, c, st) : launch_chain<MD, NTV, false, true>(ctx, c, st); \
                          return bf ? launch_chain<MD, NTV, true, false>(ctx, c, st) : launch_chain<MD, NTV, false, false>(ctx, c, st); }
    // a chain with a 64-output layer (the block's last conv) runs on the NOUT = 64 kernel; its 32-output layers issue N = 32 MMAs
#define SRB_CHW(NTV) { if (c->bwd) return bf ? launch_chain<MODE_STD, NTV, true, true, 64>(ctx, c, st) : launch_chain<MODE_STD, NTV, false, true, 64>(ctx, c, st); \
                       return bf ? launch_chain<MODE_STD, NTV, true, false, 64>(ctx, c, st) : launch_chain<MODE_STD, NTV, false, false, 64>(ctx, c, st); }
    if (c->wide) { if (c->small) SRB_CHW(2) else SRB_CHW(4) }
#undef SRB_CHW
    if (c->fold) SRB_CH(MODE_DYF, 2) else if (c->small) SRB_CH(MODE_STD, 2) else SRB_CH(MODE_STD, 4)
#undef SRB_CH
}
