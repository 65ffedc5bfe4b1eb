#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) k1(int iters, float* sink) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f, a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 0.999f, c = 1e-3f;
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
        a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
        a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
    }
    float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 123456.789f) sink[0] = s;
}
__global__ void __launch_bounds__(256) k2(int iters, float* sink) {
    float2 a0 = make_float2(threadIdx.x * 1e-3f, 1.f), a1 = make_float2(a0.x + 1.f, 2.f), a2 = make_float2(a0.x + 2.f, 3.f), a3 = make_float2(a0.x + 3.f, 4.f);
    float2 a4 = make_float2(a0.x + 4.f, 1.5f), a5 = make_float2(a0.x + 5.f, 2.5f), a6 = make_float2(a0.x + 6.f, 3.5f), a7 = make_float2(a0.x + 7.f, 4.5f);
    const float2 m = make_float2(0.999f, 0.998f), c = make_float2(1e-3f, 2e-3f);
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
        a0 = __ffma2_rn(a0, m, c); a1 = __ffma2_rn(a1, m, c); a2 = __ffma2_rn(a2, m, c); a3 = __ffma2_rn(a3, m, c);
        a4 = __ffma2_rn(a4, m, c); a5 = __ffma2_rn(a5, m, c); a6 = __ffma2_rn(a6, m, c); a7 = __ffma2_rn(a7, m, c);
    }
    float s = a0.x + a1.x + a2.x + a3.x + a4.x + a5.x + a6.x + a7.x + a0.y + a1.y + a2.y + a3.y + a4.y + a5.y + a6.y + a7.y;
    if (s == 123456.789f) sink[0] = s;
}
int main() {
    float* sink; cudaMalloc(&sink, 16);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    int iters = 1 << 16, blocks = 148 * 8;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(a); k1<<<blocks, 256>>>(iters, sink); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        printf("FFMA : %.2f TFLOP/s (%.3f ms)\n", 16.0 * iters * blocks * 256 / ms / 1e9, ms);
        cudaEventRecord(a); k2<<<blocks, 256>>>(iters, sink); cudaEventRecord(b); cudaEventSynchronize(b);
        cudaEventElapsedTime(&ms, a, b);
        printf("FFMA2: %.2f TFLOP/s (%.3f ms)\n", 32.0 * iters * blocks * 256 / ms / 1e9, ms);
    }
    return 0;
}
