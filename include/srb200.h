/* srb200.h — C ABI of libsrb200.so: hand-written sm_100a kernels for the super-resolution hot path.
 *
 * The reference (thnak/image_super_resolution) has no FFI of its own: the boundary it exposes is the
 * Python object protocol of torch.nn.Module / torch.jit.ScriptModule at four call sites
 * (rs.py:64, rs.py:94, train.py:55, train.py:56 — SURVEY §8(b)).  Each entry point below names the
 * reference code whose execution it replaces.  All pointers are plain device (or, where stated,
 * host) pointers, all sizes plain ints; no torch types.  Every function returns 0 on success or a
 * negative SRB_E* code; srb_last_error() gives the thread-local message.  No call allocates device
 * memory behind the caller's back except srb_create / srb_pack_conv / srb_model_pt_create (which
 * own the packed weights) and the *_host convenience entry (which owns pinned staging buffers).
 * All device work is enqueued on the caller's stream (a cudaStream_t passed as void*).
 *
 * Activation layout ("C8 planes"): [N][C/8][H][W][8] 16-bit floats (fp16 or bf16) — a 16-byte granule
 * holds 8 consecutive channels of one pixel.  A conv may read / write a slice of planes of a wider
 * buffer (c8_total, c8_off), which makes torch.cat of the dense blocks (utils/models.py:267-270) free.
 */
#ifndef SRB200_H
#define SRB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRB_OK 0
#define SRB_EINVAL (-1)   /* bad argument: shape, alignment, unsupported channel count ...        */
#define SRB_ECUDA (-2)    /* a CUDA runtime / driver call failed                                  */
#define SRB_ENOSUP (-3)   /* no sm_100a kernel for this configuration (never falls back to a CPU) */
#define SRB_ENOMEM (-4)

#define SRB_DT_F16 0
#define SRB_DT_BF16 1

/* epilogue flags of srb_conv2d_fwd */
#define SRB_F_LRELU 1      /* LeakyReLU(slope) after bias (in-place activations of Conv/ConvWithoutBN) */
#define SRB_F_RESIDUAL 2   /* y = out_scale*act(conv+b) + res_scale*residual                        */
#define SRB_F_TANH 4       /* tanh (tail conv, utils/models.py:636,691)                              */
#define SRB_F_U8OUT 8      /* TanhToArrayImage fused: uint8 NCHW output (utils/models.py:448-451)    */
#define SRB_F_PIXSHUF2 16  /* nn.PixelShuffle(2) folded into the store (utils/models.py:580-583)     */
#define SRB_F_F32OUT 32    /* tail conv: fp32 NCHW output                                            */
#define SRB_F_MASK 64      /* backward: output planes [mask_p0,mask_p1) *= LeakyReLU'(saved activation) */
#define SRB_F_HWC_BGR 128  /* tail conv with SRB_F_U8OUT: store NHWC, channels reversed (rs.py:65-74 glue folded into the store) */

typedef struct srb_ctx srb_ctx;
typedef struct srb_weights srb_weights;
typedef struct srb_net srb_net;
typedef void* srb_stream_t; /* cudaStream_t */

const char* srb_last_error(void);
int srb_version(void);

/* One context per device and host thread (the reference is single-threaded per process: rs.py, train.py). */
int srb_create(int device, srb_ctx** out);
void srb_destroy(srb_ctx* ctx);
int srb_num_sms(const srb_ctx* ctx);
/* Host-side options of a context (no effect inside any kernel):
 *   SRB_OPT_MAX_GRID  cap on the persistent grids (0 = one CTA per SM): leaves SMs to a co-running stream; the parity tests
 *                     use it to move the row segments' boundaries
 *   SRB_OPT_PATH      SRB_PATH_* bits selecting an alternative kernel family for a layer of the model.pt graph (the tiled
 *                     implicit-GEMM kernels the row-fold ones replaced: kept as tested second implementations) */
#define SRB_OPT_MAX_GRID 1
#define SRB_OPT_PATH 2
#define SRB_PATH_UNFUSED_TRUNK 8          /* ResidualBlock1 as two conv launches instead of the fused kernel          */
#define SRB_PATH_NO_PDL 1024              /* plain launches (no programmatic dependent launch)                         */
#define SRB_PATH_TILED_CONV1 8192         /* conv1 + long skip on the tiled kernel                                      */
#define SRB_PATH_TILED_TAIL 16384         /* conv2 9x9 + tanh on the tiled (dx-folded) kernel                           */
#define SRB_PATH_C8_HEAD 32768            /* conv0 9x9 on C8 input planes (pixel pairs in K) instead of pair granules   */
#define SRB_PATH_LARGE_TILING 65536       /* never pick the small-problem super-tile shapes                             */
#define SRB_PATH_UNSORTED_UNITS 131072    /* batched weight gradient: units in job order                                */
#define SRB_PATH_WS_LARGE_TILING 262144   /* large tiling for the weight-streaming convs only                           */
#define SRB_PATH_NO_PAIR 1048576          /* fused residual block on single CTAs (cta_group::1) instead of CTA pairs          */
#define SRB_PATH_GROUP_BOXES 2097152      /* row-fold conv1 / tail: one TMA box per 8-px group instead of whole-row boxes       */
#define SRB_PATH_PAIR4_HEAD 4194304       /* Normalize as its own pass into pixel-pair granules + conv0 on the tiled kernel    */
#define SRB_PATH_SPLIT_C1TAIL 8388608     /* conv1 + skip and conv2 9x9 + tanh as two launches instead of the fused kernel          */
int srb_set_option(srb_ctx* ctx, int key, int value);

/* Fold BN (fuse_conv_and_bn, utils/models.py:366-406: bn4 = gamma,beta,mean,var each [cout]; NULL for
 * none), convert to dtype and lay the filter out as the UMMA B operand.  w is fp32 OIHW on the HOST.
 * Replaces Conv.fuseforward / ConvWithoutBN._forward_impl weight handling (utils/models.py:97-106,191-196).
 * Supported: (cin=3,k=9) head; k=3 with cin in {32,64} (filter resident in shared memory) or cin%32==0 above 64
 * (filter streamed in 32-channel K blocks); (cout=3, cin=64, k=9) tail. */
int srb_pack_conv(srb_ctx* ctx, const float* w_oihw, const float* bias, const float* bn4, float bn_eps,
                  int cout, int cin, int k, int dtype, srb_weights** out);
void srb_free_weights(srb_weights* w);

/* One conv + fused epilogue.  Replaces aten::_convolution + bias + leaky_relu_/tanh + add/mul groups of
 * Conv/ConvWithoutBN/ResidualBlock1/RDB/RRDB/Scaler forward (utils/models.py:97-98,195-196,208-209,
 * 265-271,316-317,580-583).  x, y, residual are C8-plane device buffers (see header comment); for the
 * tail conv (cout=3) y is NCHW uint8 (SRB_F_U8OUT) or fp32 (SRB_F_F32OUT). */
int srb_conv2d_fwd(srb_ctx* ctx, const srb_weights* w, const void* x, int x_c8_total, int x_c8_off, void* y,
                   int y_c8_total, int y_c8_off, const void* residual, int res_c8_total, int res_c8_off,
                   float res_scale, float out_scale, float lrelu_slope, int flags, int N, int H, int W,
                   srb_stream_t stream);

/* Same with a second residual: y = out_scale*act(conv+b) + res_scale*residual + res2_scale*res2 — the last conv of
 * the third RDB of an RRDB carries both the RDB skip and the RRDB skip (utils/models.py:271 and :317). */
int srb_conv2d_fwd_res2(srb_ctx* ctx, const srb_weights* w, const void* x, int x_c8_total, int x_c8_off, void* y,
                        int y_c8_total, int y_c8_off, const void* residual, int res_c8_total, int res_c8_off,
                        float res_scale, const void* res2, int res2_c8_total, int res2_c8_off, float res2_scale,
                        float out_scale, float lrelu_slope, int flags, int N, int H, int W, srb_stream_t stream);

/* General form of the conv launch (the two functions above fill this struct).  The extra fields serve the
 * backward pass of train.py:57 (autograd of RDB.forward, utils/models.py:265-271): the data gradient of a conv
 * is a conv of dY with the transposed, 180-degree-rotated filter (srb_pack_conv_device, transpose=1) that
 * ACCUMULATES into a slice of the dense block's gradient buffer (residual == y, res_nplanes limits the planes that
 * carry a residual) and applies the LeakyReLU backward mask of the slice that becomes final with this launch. */
typedef struct srb_conv_args {
    const void* x; int x_c8_total, x_c8_off;
    void* y; int y_c8_total, y_c8_off;
    const void* residual; int res_c8_total, res_c8_off; int res_nplanes /* 0 = all output planes */; float res_scale;
    const void* res2; int res2_c8_total, res2_c8_off; float res2_scale;
    float out_scale, lrelu_slope;
    int flags;
    const void* mask; int mask_c8_total, mask_c8_off /* plane of the mask buffer matching output plane 0 */;
    int mask_p0, mask_p1 /* output planes [p0,p1) are masked */; float mask_slope;
    int N, H, W;
} srb_conv_args;
int srb_conv2d_ex(srb_ctx* ctx, const srb_weights* w, const srb_conv_args* args, srb_stream_t stream);

/* Conv chain: up to 4 weight-streaming 3x3 convs with 32 outputs each that feed one another through one dense buffer — conv1 ..
 * conv3 of RDB.forward (utils/models.py:267-270: o_k = act(conv_k(cat[x, o_0 .. o_{k-1}]))) or the gathered data-gradient
 * stages of its backward (train.py:57) — as ONE launch: persistent CTAs run the layers back to back and a tile of layer l
 * waits only for the 3 x 3 neighbourhood of tiles of layer l-1 (arrival counters in device memory owned by the chain).
 * args[l]: the srb_conv2d_ex arguments of layer l (same N, H, W; flags SRB_F_LRELU and / or SRB_F_MASK; no residual).
 * The buffers and packed weights named at creation must keep their addresses while the chain lives; the weights' CONTENTS
 * may be re-packed between runs.  Same results as the layers launched one by one with srb_conv2d_ex (bit-identical). */
typedef struct srb_conv_chain srb_conv_chain;
struct srb_conv_args;
int srb_conv_chain_create(srb_ctx* ctx, int n, const srb_weights* const* weights, const struct srb_conv_args* args, srb_conv_chain** out);
int srb_conv_chain_run(srb_ctx* ctx, srb_conv_chain* chain, srb_stream_t stream);
void srb_conv_chain_destroy(srb_conv_chain* chain);

/* Training keeps fp32 master weights on the DEVICE (nn.Parameter storage); these re-derive the packed 16-bit
 * operand from them every step without a host round trip.  srb_alloc_conv creates an empty packed filter for a
 * (cout, cin, k) conv; srb_pack_conv_device fills it from w_src = fp32 [src_o][src_i][k][k] on the device:
 *   transpose = 0:  W'[o][c] = w_src[o_off + o][c_off + c]            (forward operand of a cout/cin slice)
 *   transpose = 1:  W'[o][c][dy][dx] = w_src[c_off + c][o_off + o][k-1-dy][k-1-dx]   (data-gradient operand)
 * bias_src (device, may be NULL -> zero bias) is read at [o_off + o] when transpose = 0 and ignored otherwise. */
int srb_alloc_conv(srb_ctx* ctx, int cout, int cin, int k, int dtype, srb_weights** out);
/* weight_streaming != 0: a 3x3 filter with cin = 64 is laid out in 32-channel K blocks like the wider ones (streamed with the
 * halo tiles instead of resident in shared memory), which lets the conv join a srb_conv_chain. */
int srb_alloc_conv_ex(srb_ctx* ctx, int cout, int cin, int k, int dtype, int weight_streaming, srb_weights** out);
int srb_pack_conv_device(srb_ctx* ctx, srb_weights* w, const float* w_src, int src_o, int src_i, int o_off, int c_off,
                         int transpose, const float* bias_src, srb_stream_t stream);

/* Normalize.forward (utils/datasets.py:65-71, uint8 branch baked into model.pt):
 * uint8 NCHW [N,3,H,W] -> ((x/255)-mean)/std as one C8 plane (channels 3..7 zero). */
int srb_u8nchw_to_c8(srb_ctx* ctx, const uint8_t* in, void* out_c8, const float mean[3], const float std_[3],
                     float max_pixel_value, int dtype, int N, int H, int W, srb_stream_t stream);
/* Layout converters for float callers (class nets take normalised fp32 NCHW, utils/models.py:645). */
int srb_f32nchw_to_c8(srb_ctx* ctx, const float* in, void* out_c8, int C, int c8_total, int c8_off, int dtype,
                      int N, int H, int W, srb_stream_t stream);
int srb_c8_to_f32nchw(srb_ctx* ctx, const void* in_c8, float* out, int C, int c8_total, int c8_off, int dtype,
                      int N, int H, int W, srb_stream_t stream);

/* The whole model.pt graph (Model.forward, utils/models.py:738-739, on the traced old-revision Denoise):
 * Normalize -> conv0 9x9+LReLU -> 8 x ResidualBlock1 -> conv1 + long skip -> conv2 9x9 + tanh -> uint8.
 * weights: the 19 conv filters/biases in graph order (conv0, 8 x (m.0, m.1), conv1, conv2), fp32 host
 * pointers w[i] = OIHW, b[i]; mean/std/max_pixel_value are net.0.* of the archive. */
int srb_model_pt_create(srb_ctx* ctx, const float* const* w, const float* const* b, const float mean[3],
                        const float std_[3], float max_pixel_value, float lrelu_slope, int dtype, srb_net** out);
void srb_model_pt_destroy(srb_net* net);
size_t srb_model_pt_workspace_bytes(int N, int H, int W);
/* device uint8 NCHW -> device uint8 NCHW; replaces `model(window_img)` (rs.py:94) / `model(frames)` (rs.py:64) */
int srb_model_pt_forward(srb_ctx* ctx, srb_net* net, const uint8_t* in_nchw, uint8_t* out_nchw, int N, int H,
                         int W, void* workspace, size_t ws_bytes, srb_stream_t stream);
/* The video path's output glue folded into the tail conv's store: out_layout SRB_LAYOUT_NHWC_BGR
 * writes [N][H][W][3] with the channels reversed — what rs.py:65-74 builds from model(frames) with RGB2BGR
 * (utils/datasets.py:113-124) + permute + contiguous before FFMPEG_recorder.writeFrame (utils/ffmpeg.py:69-76). */
#define SRB_LAYOUT_NCHW_RGB 0
#define SRB_LAYOUT_NHWC_BGR 1
#define SRB_LAYOUT_NHWC_RGB 2      /* input only */
int srb_model_pt_forward_frames(srb_ctx* ctx, srb_net* net, const uint8_t* in_nchw, uint8_t* out, int out_layout, int N, int H,
                                int W, void* workspace, size_t ws_bytes, srb_stream_t stream);
/* The ingest half of the same glue: the frames may arrive as uint8 NCHW RGB (what the reference's
 * reader yields, utils/datasets.py:431-463, and rs.py:60-63 turns into `norm_(frames.to(device).float())`) or as uint8 HWC
 * frames in RGB or BGR order (what a decoder hands over); Normalize (utils/datasets.py:65-71) is part of conv0's loader
 * either way, so `frames.float()`, `norm_` and any permute never run as passes of their own. */
int srb_model_pt_forward_io(srb_ctx* ctx, srb_net* net, const uint8_t* in, int in_layout, uint8_t* out, int out_layout, int N,
                            int H, int W, void* workspace, size_t ws_bytes, srb_stream_t stream);
/* Normalize + conv0 (9x9, 3 -> 64, BN folded) + LeakyReLU in one launch, uint8 frames in (in_layout as above), C8 planes
 * [N][8][H][W][8] out: Normalize.forward (utils/datasets.py:65-71) fused into Conv._forward_impl of conv0
 * (utils/models.py:97-98, :687) — rs.py:62-64's `norm_(frames)` + the first layer of `model(frames)`. */
typedef struct srb_head srb_head;
int srb_head_create(srb_ctx* ctx, const float* w_64x3x9x9, const float* bias64, int dtype, srb_head** out);
void srb_head_destroy(srb_head* h);
int srb_head_fwd(srb_ctx* ctx, const srb_head* h, const uint8_t* in, int in_layout, const float mean[3], const float std_[3],
                 float max_pixel_value, void* y_c8, float lrelu_slope, int N, int H, int W, srb_stream_t stream);
/* host uint8 NCHW -> host uint8 NCHW (H2D, forward, D2H, sync) — the call a non-torch host would make */
int srb_model_pt_forward_host(srb_ctx* ctx, srb_net* net, const uint8_t* in_host, uint8_t* out_host, int N,
                              int H, int W);
/* Tile transfer of the still-image path (rs.py:92 `window_img.unsqueeze(0).to(device)`, rs.py:105-111 `result_image[...] =
 * frame.cpu()`): `planes` 2-D copies (cudaMemcpy2DAsync) of width x height bytes between a pitched host image and a device
 * tile, on the caller's stream.  kind: SRB_COPY_H2D / SRB_COPY_D2H / SRB_COPY_D2D.  Pin the host image for asynchronous copies. */
#define SRB_COPY_H2D 0
#define SRB_COPY_D2H 1
#define SRB_COPY_D2D 2
int srb_copy_rect_u8(srb_ctx* ctx, void* dst, size_t dst_pitch, size_t dst_plane, const void* src, size_t src_pitch,
                     size_t src_plane, int planes, size_t width, size_t height, int kind, srb_stream_t stream);
/* number of kernels of this library one srb_model_pt_forward enqueues for H x W images (10: head with the ingest folded
 * in, 8 fused residual blocks, conv1 + tail; 11 for rows under 57 pixels: conv1 and tail as two launches) */
int srb_model_pt_launches(int H, int W);
/* measurement only: record a CUDA event before/after every launch of srb_model_pt_forward on its stream;
 * srb_profile_read returns the per-launch milliseconds of the last forward (count written, <0 on error) */
/* cudaDeviceSynchronize + report a kernel-side pipeline timeout (kernels never hang: bounded waits) */
int srb_check_device(srb_ctx* ctx);
/* The same report without a synchronisation: reads the host-mapped mirror of the kernels' abort flag.  SRB_OK when no
 * wait of any launch on this device has timed out since the last check; otherwise SRB_ECUDA (flag cleared) — every result
 * produced since is invalid.  The Python layer calls it wherever it hands results to its caller. */
int srb_poll_abort(srb_ctx* ctx);
/* on != 0: every conv launch of this context carries cudaLaunchAttributeProgrammaticStreamSerialization, so its prologue
 * (barriers, TMEM, filters, bias) overlaps the tail of the previous launch in the stream (also as programmatic edges of a
 * captured CUDA graph).  The kernels wait for the previous launch before touching activations, but they read packed weights
 * and biases in the prologue: only enable this where those are not written by the immediately preceding launches of the
 * same stream.  srb_model_pt_forward / srb_resblock_fwd / srb_rowconv_fwd / srb_tail9_fwd always launch this way. */
int srb_set_programmatic_launch(srb_ctx* ctx, int on);
/* Timing of srb_model_pt_forward's launches with CUDA events on the launching stream.  on = 1: an event after every launch (an
 * event between two launches ends their programmatic overlap: durations of the kernels in isolation); on = 2: events only before
 * the first and after the last of the trunk's eight launches, which overlap through programmatic dependent launch (srb_profile_read
 * then returns three intervals: input pack + head conv, the trunk, conv1 + tail) — the trunk interval / 8 is a fused block's average
 * duration inside the pipeline; 0: off. */
int srb_profile_enable(srb_ctx* ctx, int on);
int srb_profile_read(srb_ctx* ctx, float* ms, int max_n);
/* (probe hooks — in-kernel bottleneck probes, clock traces, the tcgen05.mma micro-benchmark — are not part of this library:
 * they exist only in the -DSRB_DEBUG build, libsrb200_dbg.so, declared in include/srb200_debug.h and loaded by tools/) */

/* Fused residual block  y = x + convB(LeakyReLU(convA(x)))  — ResidualBlock1.forward (utils/models.py:208-209) with
 * m = Sequential(Conv(64,32,3,act=LeakyReLU), Conv(32,64,3,act=False)), BN already folded (model.pt state: net.1.residual.{i}.m.{0,1}).
 * One kernel: the 32-channel hidden activation never leaves shared memory.  wA [32][64][3][3], wB [64][32][3][3] fp32 host.
 * x, y: 8-plane C8 buffers (distinct).  Every H, W > 0 is covered (srb_resblock_supported is kept for callers that asked
 * before every width was): rows up to 128 px are tiled several images to an M-tile, wider rows as overlapping strips. */
typedef struct srb_resblock srb_resblock;
int srb_resblock_create(srb_ctx* ctx, const float* wA, const float* bA, const float* wB, const float* bB, int dtype,
                        srb_resblock** out);
void srb_resblock_destroy(srb_resblock* rb);
int srb_resblock_supported(int H, int W);
int srb_resblock_fwd(srb_ctx* ctx, const srb_resblock* rb, const void* x, void* y, float lrelu_slope, int N, int H, int W,
                     srb_stream_t stream);

/* Head conv on pixel-pair granules (model.pt: Normalize + net.1.conv0, utils/datasets.py:65-71, utils/models.py:97-98).
 * srb_u8nchw_to_pair4 writes, per image row, W + 1 granules [pixel g | pixel g+1] (4 x 16-bit each: 3 normalised channels + 0;
 * g = -1 .. W-1, out-of-image pixels zero) — `out` holds N*H*(W+1) granules.  srb_pack_conv_pair4 packs the [64][3][9][9]
 * filter so that one K=16 MMA step covers four dx taps (27 steps instead of 45); the result is used with srb_conv2d_fwd
 * (x_c8_total = 1) exactly like the srb_pack_conv head conv, and gives the same values. */
int srb_u8nchw_to_pair4(srb_ctx* ctx, const uint8_t* in_nchw, void* out_pairs, const float mean[3], const float std_[3],
                        float max_pixel_value, int dtype, int N, int H, int W, srb_stream_t stream);
int srb_pack_conv_pair4(srb_ctx* ctx, const float* w_oihw, const float* bias, int cout, int dtype, srb_weights** out);

/* Row-fold 3x3 conv 64 -> 64: y = act(conv(x) + b) + res — `Conv(64,64,3,act=False)` utils/models.py:690 with the long
 * skip `inputs + self.conv1(residual)` :704 (model.pt: net.1.conv1), the same GEMM mapping as the fused residual block
 * (dy taps folded into N = 192, accumulation over adjacent TMEM rows).  w [64][64][3][3] fp32 host (BN folded),
 * res: 8-plane C8 buffer or NULL, lrelu_slope applied when flags has SRB_F_LRELU.  Any H, W > 0. */
typedef struct srb_rowconv srb_rowconv;
int srb_rowconv_create(srb_ctx* ctx, const float* w, const float* bias, int dtype, srb_rowconv** out);
void srb_rowconv_destroy(srb_rowconv* rc);
int srb_rowconv_fwd(srb_ctx* ctx, const srb_rowconv* rc, const void* x, const void* res, void* y, float lrelu_slope, int flags,
                    int N, int H, int W, srb_stream_t stream);

/* Row-fold 9x9 conv 64 -> 3 + bias (+ tanh) -> NCHW uint8 / fp32: `ConvWithoutBN(64,3,9,act=Tanh)` utils/models.py:691 and
 * `TanhToArrayImage.forward` :448-451 (model.pt: net.1.conv2, net.2) — both tap directions folded into N, accumulation over
 * 9 adjacent TMEM rows.  w [3][64][9][9] fp32 host; x: 8-plane C8 buffer; y: NCHW uint8 (SRB_F_U8OUT) or fp32 (SRB_F_F32OUT);
 * flags: SRB_F_TANH | one of the output formats.  Any H, W > 0. */
typedef struct srb_tail9 srb_tail9;
int srb_tail9_create(srb_ctx* ctx, const float* w, const float* bias, int dtype, srb_tail9** out);
void srb_tail9_destroy(srb_tail9* t);
int srb_tail9_fwd(srb_ctx* ctx, const srb_tail9* t, const void* x, void* y, int flags, int N, int H, int W, srb_stream_t stream);
/* conv1 + long skip + conv2 + tanh + TanhToArrayImage as ONE launch (utils/models.py:690, :704, :691, :448-451): srb_rowconv_fwd
 * followed by srb_tail9_fwd with conv1's 64-channel result kept in shared memory instead of HBM (bit-identical to the two calls).
 * Takes the packed filters of an srb_rowconv and an srb_tail9 of the same operand type; rows of at least 57 pixels
 * (srb_c1tail_supported), otherwise make the two calls.  flags as srb_tail9_fwd. */
int srb_c1tail_supported(int H, int W);
int srb_c1tail_fwd(srb_ctx* ctx, const srb_rowconv* rc, const srb_tail9* t, const void* x, const void* skip, void* y, int flags,
                   int N, int H, int W, srb_stream_t stream);

/* Weight gradient of a 3x3 conv (autograd of aten::_convolution, train.py:57):
 * dw[co][ci][dy][dx] += scale * sum_p dY[p][co] * X[p+(dy-1,dx-1)][ci].  x, dy: C8-plane slices (device);
 * dw: fp32 OIHW [cout][cin][3][3] on the device, ACCUMULATED into (zero it first).  fp32 atomics: summation order
 * across CTAs is not fixed.  Bias gradient: db[c] += scale * sum_p dY[p][c]. */
int srb_conv2d_wgrad(srb_ctx* ctx, const void* x, int x_c8_total, int x_c8_off, int cin, const void* dy, int dy_c8_total,
                     int dy_c8_off, int cout, float* dw_oihw, float scale, int k, int dtype, int N, int H, int W,
                     srb_stream_t stream);
int srb_bias_grad(srb_ctx* ctx, const void* dy, int dy_c8_total, int dy_c8_off, int cout, float* db, float scale, int dtype,
                  int N, int H, int W, srb_stream_t stream);

/* 9x9 weight gradients of the head (cin=3,cout=64; utils/models.py:625) and tail (cin=64,cout=3; :636) convs on the same
 * tensor-core GEMM (9 row shifts of one dY plane folded into M, 9 dx accumulators; any cin, cout).
 * Same contract as srb_conv2d_wgrad: dw fp32 OIHW [cout][cin][9][9], accumulated into. */
int srb_conv9_wgrad(srb_ctx* ctx, const void* x, int x_c8_total, int x_c8_off, int cin, const void* dy, int dy_c8_total,
                    int dy_c8_off, int cout, float* dw_oihw, float scale, int dtype, int N, int H, int W, srb_stream_t stream);

/* Backward glue that is not a conv epilogue (autograd of EResNet.forward, utils/models.py:645-650):
 * tanh' of the tail conv: g_c8[n,y,x,0..2] = scale * gout * (1 - out^2), gout/out fp32 NCHW [N,3,H,W], g_c8 one C8 plane. */
int srb_tanh_bwd_to_c8(srb_ctx* ctx, const float* gout, const float* out, void* g_c8, float scale, int dtype, int N, int H, int W,
                       srb_stream_t stream);
/* Scaler backward (utils/models.py:580-583): g_s[n,4c+2i+j,y,x] = g_u[n,c,2y+i,2x+j] * LeakyReLU'(u); g_u, u: c8_u planes at
 * 2H x 2W; g_s: 4*c8_u planes at H x W. */
int srb_pixunshuffle_mask(srb_ctx* ctx, const void* g_u, const void* u, void* g_s, int c8_u, float slope, int dtype, int N, int H,
                          int W, srb_stream_t stream);
/* Bare nn.PixelShuffle(2) + LeakyReLU of the current-revision Denoise generator (utils/models.py:685 `residual_conv1`; the conv
 * before it adds a residual first, so its store cannot carry the shuffle): u[n, c, 2y+i, 2x+j] = lrelu(s[n, 4c+2i+j, y, x]).
 * s: 4*c8_u planes at H x W, u: c8_u planes at 2H x 2W.  With slope = 1 and the roles swapped, srb_pixunshuffle_mask (u = g_u,
 * slope 1) is the inverse permutation: the space-to-depth step that turns the stride-2 conv (:680) into a stride-1 one. */
int srb_pixshuffle_lrelu(srb_ctx* ctx, const void* s_c8, void* u_c8, int c8_u, float slope, int dtype, int N, int H, int W,
                         srb_stream_t stream);
/* out = (alpha*a + beta*b) * LeakyReLU'(act) over C8 plane slices; b and act may be NULL (gradient fan-in at the long skip). */
int srb_c8_axpby_mask(srb_ctx* ctx, const void* a, int a_c8_total, int a_c8_off, float alpha, const void* b, int b_c8_total,
                      int b_c8_off, float beta, const void* act, int act_c8_total, int act_c8_off, float slope, void* out,
                      int out_c8_total, int out_c8_off, int planes, int dtype, int N, int H, int W, srb_stream_t stream);

/* Fused optimizer tail over FLAT fp32 arenas (the parameters / gradients / Adam moments / EMA weights of the whole net as
 * one zero-padded buffer each, length n a multiple of 4) — SURVEY §8(f) rank 1.
 * srb_grad_stats replaces the per-tensor loops of gradscaler.unscale_ (train.py:58: inf/nan check of g/scale) and of
 * clip_grad_norm_ (train.py:59: global 2-norm): stats[0] = sum (g/scale)^2, stats[1] = 1 if any g/scale is non-finite,
 * stats[2] is left alone (the number of optimizer steps taken: srb_adam_ema_step advances it; zero it once), so the
 * statistics may be recomputed any number of times per step.  scale_dev = GradScaler's device scalar or NULL.
 * workspace: srb_optim_workspace_floats() floats, zeroed once.  Deterministic (fixed-order partial sums).
 * srb_adam_ema_step replaces gradscaler.step(Adam) (train.py:60; torch.optim.Adam arithmetic, no amsgrad; skipped on the
 * device when stats[1] != 0) and ema.update (train.py:63 -> utils/models.py:31-40: ema = d*ema + (1-d)*p, ema may be
 * NULL) in one pass; g is read as g / scale * min(1, max_norm / (sqrt(stats[0]) + 1e-6)) (max_norm <= 0: no clipping)
 * and never rewritten.  The step number of the update is stats[2] + 1; the kernel's last block stores it back (not when
 * the step is skipped).  workspace: the same buffer srb_grad_stats uses (its arrival counter is shared). */
int srb_optim_workspace_floats(void);
int srb_grad_stats(srb_ctx* ctx, const float* grads, size_t n, const float* scale_dev, float* stats, float* workspace,
                   srb_stream_t stream);
int srb_adam_ema_step(srb_ctx* ctx, float* params, const float* grads, float* exp_avg, float* exp_avg_sq, float* ema, size_t n,
                      const float* scale_dev, float* stats, float* workspace, float max_norm, float lr, float beta1, float beta2,
                      float eps, float weight_decay, float ema_decay, srb_stream_t stream);

/* Batched gradients: every 3x3 weight gradient and every bias gradient of one backward pass as ONE launch each.
 * Weight gradients are off backpropagation's critical path, so the step defers them to its end; a unit
 * (conv, <= 160-ch cin block, 32-ch cout block) is then owned by one CTA (splitk = 1): no cross-CTA reduction.  The bias
 * gradient of a conv that also has a weight gradient rides on its first cin block (the dY tiles are in shared memory).
 * Buffers named by the jobs must stay allocated (and keep their addresses) while the batch object lives.
 * dw / db are ACCUMULATED into; either may be NULL. */
typedef struct srb_grad_job {
    const void* x; int x_c8_total, x_c8_off, cin;
    const void* dy; int dy_c8_total, dy_c8_off, cout;
    float* dw; float* db; float scale;
    int N, H, W;
} srb_grad_job;
typedef struct srb_grad_batch srb_grad_batch;
int srb_grad_batch_create(srb_ctx* ctx, const srb_grad_job* jobs, int n, int dtype, srb_grad_batch** out);
int srb_grad_batch_units(const srb_grad_batch* b);
int srb_grad_batch_run(srb_ctx* ctx, srb_grad_batch* b, int splitk, srb_stream_t stream);
void srb_grad_batch_destroy(srb_grad_batch* b);

/* Batched srb_pack_conv_device: all packed operands of a net re-derived from the fp32 masters in one launch.
 * c_n > 0: the job fills only the operand's input channels [c_lo, c_lo + c_n) from source channels c_off .. (3x3 operands):
 * an operand whose input channels are gathered from several source filters takes one job per source — the data-gradient
 * convs of a dense block read the concatenated output gradients of all later convs (training.py).  scale (0 = 1) multiplies
 * the source weights. */
typedef struct srb_pack_job {
    srb_weights* w; const float* w_src; int src_o, src_i, o_off, c_off, transpose; const float* bias_src;
    int c_lo, c_n; float scale;
} srb_pack_job;
typedef struct srb_pack_batch srb_pack_batch;
int srb_pack_batch_create(srb_ctx* ctx, const srb_pack_job* jobs, int n, srb_pack_batch** out);
int srb_pack_batch_run(srb_ctx* ctx, srb_pack_batch* b, srb_stream_t stream);
void srb_pack_batch_destroy(srb_pack_batch* b);

/* L1Loss.forward / backward (utils/loss.py:38-39): loss = w * mean|pred - tgt| ; dpred = w*scale*sign/numel.
 * pred_dtype: SRB_DT_F16 / SRB_DT_BF16 / 2 = fp32.  tgt fp32.  partials: >= srb_l1_partials_len() floats. */
int srb_l1_partials_len(void);
int srb_l1_loss_fwd(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                    float* loss_out, float* partials, srb_stream_t stream);
int srb_l1_loss_bwd(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                    float grad_scale, void* dpred, srb_stream_t stream);

/* Allocation-free variant of srb_l1_loss_fwd.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The reduction scratch that srb_l1_loss_fwd takes from the caller (`partials`) lives in a file-scope
 * `static __device__ __align__(256)` array of the library (16.6 KB, held by every GPU that loads the library until exit);
 * the call performs no cudaMalloc / cudaFree / allocating routine.  Limit: n <= SRB_L1_NOALLOC_MAX_ELEMS (4096 partial sums
 * x 16 384 elements each) — larger inputs return SRB_EINVAL.  All calls on a device share that array: calls that overlap
 * on the GPU (different streams) race; calls on one stream are safe (the arrival counter is cleared on the stream before
 * every launch, the partial sums are rewritten).  Same value as srb_l1_loss_fwd up to the summation order (rel 1e-6). */
#define SRB_L1_NOALLOC_MAX_ELEMS 67108864
int srb_l1_loss_fwd_noalloc(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                            float* loss_out, srb_stream_t stream);

/* same, times a device scalar *grad_out_dev (the upstream gradient of the loss, e.g. GradScaler's scale: train.py:57) */
int srb_l1_loss_bwd_dev(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                        float grad_scale, const float* grad_out_dev, void* dpred, srb_stream_t stream);

/* Train-mode nn.BatchNorm2d of the reference's `Conv` block (utils/models.py:75-111: conv -> BatchNorm2d -> act), for the `ResNet`
 * generator trained by train.py:41-67.  The conv stores its raw output y (16-bit C8 planes, no bias, no
 * activation); the four passes below work on a slice of `planes` planes (8 channels each) of the batch N x H x W:
 *   srb_bn_stats        sums[c] += sum y, sums[C + c] += sum y^2           (the caller zeroes `sums`, 2 * C doubles, before the step)
 *   srb_bn_apply        out = out_scale * act(gamma * (y - mean) / sqrt(var + eps) + beta) + res_scale * res + res2_scale * res2
 *                       with the biased batch variance; moves running_mean / running_var (momentum, unbiased variance) and
 *                       num_batches_tracked exactly as torch does in training mode (all three may be NULL)
 *   srb_bn_bwd_reduce   dz = up_scale * da * act'(z);  bsums[c] += sum dz, bsums[C + c] += sum dz * x_hat   (caller zeroes `bsums`)
 *   srb_bn_bwd_apply    out = dL/dy = gamma / sqrt(var + eps) * (dz - mean(dz) - x_hat * mean(dz * x_hat));
 *                       dgamma += grad_scale * sum dz * x_hat, dbeta += grad_scale * sum dz   (fp32, either may be NULL)
 * act: 0 identity, 1 LeakyReLU(slope).  Sums are accumulated in double with atomics: results are deterministic to fp32 rounding. */
typedef struct srb_bn_args {
    const void* y; int y_c8_total, y_c8_off;
    void* out; int out_c8_total, out_c8_off;
    int planes;
    const float* gamma; const float* beta;
    double* sums;
    float* running_mean; float* running_var; long long* num_batches_tracked;
    float momentum, eps;
    float slope; int act;
    float out_scale;
    const void* res; int res_c8_total, res_c8_off; float res_scale;
    const void* res2; int res2_c8_total, res2_c8_off; float res2_scale;
    const void* da; int da_c8_total, da_c8_off; float up_scale;
    double* bsums;
    float* dgamma; float* dbeta; float grad_scale;
    int N, H, W;
} srb_bn_args;
int srb_bn_stats(srb_ctx* ctx, const srb_bn_args* args, int dtype, srb_stream_t stream);
int srb_bn_apply(srb_ctx* ctx, const srb_bn_args* args, int dtype, srb_stream_t stream);
int srb_bn_bwd_reduce(srb_ctx* ctx, const srb_bn_args* args, int dtype, srb_stream_t stream);
int srb_bn_bwd_apply(srb_ctx* ctx, const srb_bn_args* args, int dtype, srb_stream_t stream);

/* nn.MSELoss() (train.py:13, :233, :264 — the loss of `train.py` without --enchant): loss = w * mean((pred - tgt)^2), and
 * dpred = w * grad_scale * (*grad_out_dev) * 2 (pred - tgt) / n.  Same kernels, scratch and determinism as the L1 pair. */
int srb_mse_loss_fwd(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                     float* loss_out, float* partials, srb_stream_t stream);
int srb_mse_loss_bwd_dev(srb_ctx* ctx, const void* pred, int pred_dtype, const float* tgt, size_t n, float w,
                         float grad_scale, const float* grad_out_dev, void* dpred, srb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SRB200_H */
