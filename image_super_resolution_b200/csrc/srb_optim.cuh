// Fused optimizer tail of the training step over FLAT fp32 arenas (parameters, gradients, Adam moments, EMA weights of the
// whole net as one buffer each) — SURVEY §8(f) rank 1.  Replaces, for the reference, the per-tensor loops of
//   gradscaler.unscale_(optimizer)                 train.py:58   (inf/nan check + 1/scale)
//   clip_grad_norm_(model.parameters(), 10)        train.py:59   (global 2-norm, coef = min(1, max_norm / (norm + 1e-6)))
//   gradscaler.step(optimizer)  [Adam]             train.py:60   (skipped when a gradient is non-finite)
//   ema.update(model)                              train.py:63 -> utils/models.py:31-40   (v = d*v + (1-d)*p)
// with two launches: grad_stats_kernel (one read of the gradients: sum of squares of the unscaled gradient, non-finite
// flag, step counter) and adam_ema_kernel (one pass over p, g, m, v, ema).  The gradients are never rewritten: the unscale
// factor and the clip coefficient are applied where the Adam moments read them.  HBM-bound, 16-byte accesses.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace srb {

// stats[0] = sum (g / scale)^2, stats[1] = 1 if any g / scale is non-finite, stats[2] = number of optimizer steps taken
// (advanced by adam_ema_kernel's last block, so this kernel may run any number of times per step: logging the norm before
// unscale_, after step(), or without a step at all leaves the bias correction alone)
constexpr int GSTAT_MAX_BLOCKS = 1024;

__global__ void __launch_bounds__(256) grad_stats_kernel(const float4* __restrict__ g, size_t n4, const float* __restrict__ scale,
                                                         float* __restrict__ stats, float* __restrict__ partials,
                                                         unsigned int* __restrict__ counter) {
    const float inv = scale != nullptr ? 1.0f / __ldg(scale) : 1.0f;
    float ss = 0.f;
    int bad = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 q = __ldg(g + i);
        const float a = q.x * inv, b = q.y * inv, c = q.z * inv, d = q.w * inv;
        ss += a * a + b * b + c * c + d * d;
        bad |= !(isfinite(a) && isfinite(b) && isfinite(c) && isfinite(d));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    bad = __any_sync(0xffffffffu, bad);
    __shared__ float s_ss[8];
    __shared__ int s_bad[8];
    __shared__ bool s_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_ss[warp] = ss; s_bad[warp] = bad; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f; int bb = 0;
        for (int w = 0; w < 8; ++w) { t += s_ss[w]; bb |= s_bad[w]; }
        partials[2 * blockIdx.x] = t;
        partials[2 * blockIdx.x + 1] = bb ? 1.f : 0.f;
        __threadfence();
        s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {          // the last block to arrive adds the per-block partials in a fixed order (deterministic norm)
        __threadfence();
        float t = 0.f, bb = 0.f;
        for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {
            t += __ldcg(partials + 2 * i);
            bb = fmaxf(bb, __ldcg(partials + 2 * i + 1));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { t += __shfl_xor_sync(0xffffffffu, t, o); bb = fmaxf(bb, __shfl_xor_sync(0xffffffffu, bb, o)); }
        __syncthreads();
        if (lane == 0) { s_ss[warp] = t; s_bad[warp] = bb > 0.f; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float tt = 0.f; int b2 = 0;
            for (int w = 0; w < 8; ++w) { tt += s_ss[w]; b2 |= s_bad[w]; }
            const bool nonfinite = b2 || !isfinite(tt);
            stats[0] = tt;
            stats[1] = nonfinite ? 1.f : 0.f;
            *counter = 0u;
        }
    }
}

struct AdamArgs {
    float4* p; const float4* g; float4* m; float4* v; float4* ema;      // ema may be nullptr
    size_t n4;
    const float* scale;      // GradScaler scale (device) or nullptr
    float* stats;            // grad_stats_kernel output of this step; [2] = steps taken so far (advanced here)
    unsigned int* counter;   // arrival counter (zero on entry, left at zero)
    float max_norm;          // > 0: clip the global norm (clip_grad_norm_ was called this step)
    float lr, beta1, beta2, eps, weight_decay;
    float ema_decay;         // d of utils/models.py:34 for this update
};

__global__ void __launch_bounds__(256) adam_ema_kernel(const AdamArgs a) {
    const float inv = a.scale != nullptr ? 1.0f / __ldg(a.scale) : 1.0f;
    // (plain loads: the last block of this launch writes stats[2] — after every block has read it, see the tail)
    const float sumsq = a.stats[0], found_inf = a.stats[1], step = a.stats[2] + 1.0f;
    const bool skip = found_inf != 0.f;                     // GradScaler.step: no optimizer step on non-finite gradients
    float coef = 1.0f;
    if (a.max_norm > 0.f) coef = fminf(a.max_norm / (sqrtf(sumsq) + 1e-6f), 1.0f);     // torch.nn.utils.clip_grad_norm_
    const float gmul = inv * coef;
    const float bc1 = 1.0f - powf(a.beta1, step), bc2 = 1.0f - powf(a.beta2, step);
    const float step_size = a.lr / bc1, rsq_bc2 = rsqrtf(bc2);
    const float d = a.ema_decay, omd = 1.0f - a.ema_decay;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < a.n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 p = a.p[i];
        if (!skip) {
            const float4 gq = __ldg(a.g + i);
            float4 m = a.m[i], v = a.v[i];
            float gg[4] = {gq.x * gmul, gq.y * gmul, gq.z * gmul, gq.w * gmul};
            float pp[4] = {p.x, p.y, p.z, p.w}, mm[4] = {m.x, m.y, m.z, m.w}, vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float g1 = gg[k] + a.weight_decay * pp[k];
                mm[k] = mm[k] + (g1 - mm[k]) * (1.0f - a.beta1);                  // exp_avg.lerp_(grad, 1 - beta1)
                vv[k] = vv[k] * a.beta2 + (1.0f - a.beta2) * g1 * g1;
                const float denom = sqrtf(vv[k]) * rsq_bc2 + a.eps;
                pp[k] -= step_size * (mm[k] / denom);
            }
            p = make_float4(pp[0], pp[1], pp[2], pp[3]);
            a.p[i] = p;
            a.m[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
            a.v[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
        }
        if (a.ema != nullptr) {
            float4 e = a.ema[i];
            e.x = d * e.x + omd * p.x; e.y = d * e.y + omd * p.y; e.z = d * e.z + omd * p.z; e.w = d * e.w + omd * p.w;
            a.ema[i] = e;
        }
    }
    // the step counter advances once per optimizer step taken: by the last block to finish (every block has read it by then)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(a.counter, 1u) == gridDim.x - 1) {
            if (!skip) a.stats[2] = step;
            *a.counter = 0u;
        }
    }
}

}  // namespace srb
