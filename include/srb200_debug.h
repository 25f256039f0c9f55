/* srb200_debug.h — probe hooks of the -DSRB_DEBUG build (libsrb200_dbg.so).  NOT part of the product library:
 * libsrb200.so exports none of these and its kernels carry none of the run-time probe branches.  Used by tools/ only. */
#ifndef SRB200_DEBUG_H
#define SRB200_DEBUG_H
#include "srb200.h"
#ifdef __cplusplus
extern "C" {
#endif
/* key 0 = swap LBO/SBO in UMMA descriptors, 1 = grid cap, 2 = probe bits (1 no TMA, 2 no MMA, 4 no global I/O, ... and the
 * SRB_PATH_* bits), 3 = extra launches of the fused block per call, 4 / 5 = block size / shared memory of the launch probe */
int srb_debug_set(srb_ctx* ctx, int key, int value);
int srb_debug_trace(srb_ctx* ctx, void* dev_buf_int64);  /* clock stamps of CTA 0 per row and per-CTA start / end times */
/* cycles for `iters` rounds of `racc` interleaved M=128 x n x 16 tcgen05.mma from `m64` issuing warps;
 * out2_host[0] = issue cycles, out2_host[1] = cycles until all MMAs completed */
int srb_ubench_mma(srb_ctx* ctx, int n, int layout, int racc, int iters, int m64, int grid, long long* out2_host);
/* Mixed issue streams: warp w issues, per "row", cnt[w] chained MMAs of N = n[w] round-robin over acc[w] accumulators; `rows`
 * rows back to back.  Reproduces the fused block's two (or three) issuers.  out_host[0] = cycles until everything completed,
 * out_host[1 + w] = warp w's issue cycles. */
int srb_ubench_mix(srb_ctx* ctx, int nwarps, const int* n, const int* cnt, const int* acc, int rows, int cta_pair, long long* out_host);
#ifdef __cplusplus
}
#endif
#endif
