// qtts_microbench.cu — instrumentation only: cost of one grid-wide barrier on this device, for the
// barrier variants considered in DESIGN.md.  Not on the product path.
#include "qtts_engine.h"

namespace qtts {

// variant 0: __syncthreads + thread 0 {__threadfence; atomicAdd; poll ld.acquire.gpu}
// variant 1: same without the __threadfence (lower bound of the counter protocol)
// variant 2: red.release.gpu + poll ld.acquire.gpu
// variant 3: variant 0 + 64 global stores per CTA before the barrier (store drain included)
// variant 4: per-CTA flags: every CTA writes flag[cta] = epoch; warp 0 polls all flags
__global__ void __launch_bounds__(kThreads, 1)
k_barrier_bench(unsigned int * count, unsigned int * flags, float * sink, int iters, int variant) {
    unsigned int epoch = 0;
    for (int it = 0; it < iters; ++it) {
        if (variant == 3 && threadIdx.x < 64) sink[blockIdx.x * 64 + threadIdx.x] = (float) it;
        __syncthreads();
        if (variant == 4) {
            if (threadIdx.x == 0) {
                __threadfence();
                asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(flags + blockIdx.x), "r"((unsigned) (it + 1)) : "memory");
            }
            if (threadIdx.x < 32) {
                for (int c = threadIdx.x; c < (int) gridDim.x; c += 32) {
                    unsigned int v;
                    do {
                        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + c) : "memory");
                    } while ((int) (v - (unsigned) (it + 1)) < 0);
                }
                __syncwarp();
                __threadfence();
            }
        } else if (threadIdx.x == 0) {
            epoch += gridDim.x;
            if (variant == 2) {
                asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(count) : "memory");
            } else {
                if (variant != 1) __threadfence();
                atomicAdd(count, 1u);
            }
            unsigned int v;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(count) : "memory");
            } while ((int) (v - epoch) < 0);
        }
        __syncthreads();
    }
}

}  // namespace qtts

extern "C" double qtts_debug_barrier_us(int device, int variant, int iters) {
    using namespace qtts;
    if (cudaSetDevice(device) != cudaSuccess) return -1.0;
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    unsigned int * count = nullptr, * flags = nullptr;
    float * sink = nullptr;
    cudaMalloc(&count, 128); cudaMalloc(&flags, 4096); cudaMalloc(&sink, 148 * 64 * 4 * 2);
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    float ms = -1.f;
    for (int rep = 0; rep < 2; ++rep) {
        cudaMemset(count, 0, 128); cudaMemset(flags, 0, 4096);
        void * args[] = {&count, &flags, &sink, &iters, &variant};
        cudaEventRecord(a);
        if (cudaLaunchCooperativeKernel((const void *) k_barrier_bench, dim3(prop.multiProcessorCount), dim3(kThreads),
                                        args, 200 * 1024 > 0 ? 0 : 0, 0) != cudaSuccess) { ms = -1.f; break; }
        cudaEventRecord(b);
        if (cudaEventSynchronize(b) != cudaSuccess) { ms = -1.f; break; }
        cudaEventElapsedTime(&ms, a, b);
    }
    cudaFree(count); cudaFree(flags); cudaFree(sink);
    cudaEventDestroy(a); cudaEventDestroy(b);
    return ms < 0 ? -1.0 : (double) ms * 1e3 / iters;
}
