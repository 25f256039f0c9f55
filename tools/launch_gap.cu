// Launch-to-launch period of back-to-back kernels as a function of the CTA footprint (shared memory, TMEM, threads).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/launch_gap tools/launch_gap.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(672, 1) k(int mode, long long spin, int* sink) {
    extern __shared__ uint8_t smem[];
    __shared__ uint32_t s_t;
    if (mode & 1) {
        if (threadIdx.x < 32) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_t)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        __syncthreads();
    }
    if (mode & 4) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (mode & 4) asm volatile("griddepcontrol.wait;" ::: "memory");
    long long t0 = clock64();
    while (clock64() - t0 < spin) { }
    if (mode & 2) smem[threadIdx.x] = (uint8_t)spin;
    __syncthreads();
    if (mode & 1) {
        if (threadIdx.x < 32) {
            uint32_t t = s_t;
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(t) : "memory");
        }
    }
    if (sink && threadIdx.x == 9999) *sink = smem[0];
}
int main() {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    for (int threads : {128, 672})
        for (int smem : {0, 100 * 1024, 220 * 1024})
            for (int mode : {0, 1, 4, 5}) {
                for (long long spin : {0LL, 40000LL}) {
                    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
                    auto launch = [&]() {
                        cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3(148); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem;
                        cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                        at[0].val.programmaticStreamSerializationAllowed = (mode & 4) ? 1 : 0; cfg.attrs = at; cfg.numAttrs = 1;
                        cudaLaunchKernelEx(&cfg, k, mode, spin, (int*)nullptr);
                    };
                    for (int i = 0; i < 10; ++i) launch();
                    cudaEventRecord(a);
                    for (int i = 0; i < 200; ++i) launch();
                    cudaEventRecord(b); cudaEventSynchronize(b);
                    float ms; cudaEventElapsedTime(&ms, a, b);
                    printf("threads %3d smem %3d KB tmem %d pdl %d spin %5lld clk: %.2f us per launch\n", threads, smem / 1024, mode & 1, (mode >> 2) & 1, spin, ms * 1000 / 200);
                }
            }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
