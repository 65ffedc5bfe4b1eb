// Back-to-back barrier cost on one SM (sm_100a): what a team barrier of kernel (b) costs when every warp arrives at
// once.  Variants: bar.sync over the whole CTA / a named sub-barrier, CTA sizes 64..384 threads, and an mbarrier
// arrive + try_wait loop (the alternative for a split arrive/wait).
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o barrier_latency barrier_latency.cu && ./barrier_latency
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void bar_named(int iters, long long* out) {
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
        asm volatile("bar.sync 1, %0;" ::"r"(blockDim.x) : "memory");
        asm volatile("bar.sync 1, %0;" ::"r"(blockDim.x) : "memory");
        asm volatile("bar.sync 1, %0;" ::"r"(blockDim.x) : "memory");
        asm volatile("bar.sync 1, %0;" ::"r"(blockDim.x) : "memory");
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) *out = t1 - t0;
}

__global__ void bar_cta(int iters, long long* out) {
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
        __syncthreads(); __syncthreads(); __syncthreads(); __syncthreads();
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) *out = t1 - t0;
}

__global__ void bar_mbarrier(int iters, long long* out) {
    __shared__ __align__(8) uint64_t mbar;
    const unsigned addr = (unsigned)__cvta_generic_to_shared(&mbar);
    if (threadIdx.x == 0) asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(addr), "r"(blockDim.x / 32));
    __syncthreads();
    const long long t0 = clock64();
    unsigned phase = 0;
#pragma unroll 1
    for (int i = 0; i < 4 * iters; ++i) {
        if ((threadIdx.x & 31) == 0) asm volatile("mbarrier.arrive.shared.b64 _, [%0];" ::"r"(addr) : "memory");
        unsigned done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(addr), "r"(phase) : "memory");
        phase ^= 1;
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0) *out = t1 - t0;
}

int main() {
    long long* d; cudaMalloc(&d, 8);
    const int iters = 5000;
    for (int threads : {32, 64, 96, 192, 384, 768}) {
        long long a = 0, b = 0, c = 0;
        bar_named<<<1, threads>>>(iters, d); cudaMemcpy(&a, d, 8, cudaMemcpyDeviceToHost);
        bar_cta<<<1, threads>>>(iters, d); cudaMemcpy(&b, d, 8, cudaMemcpyDeviceToHost);
        bar_mbarrier<<<1, threads>>>(iters, d); cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
        printf("%3d threads: bar.sync named %.1f  __syncthreads %.1f  mbarrier arrive+try_wait %.1f cycles\n", threads,
               a / (4.0 * iters), b / (4.0 * iters), c / (4.0 * iters));
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
