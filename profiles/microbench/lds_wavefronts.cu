// Shared-memory load cost by access width and lane->address pattern on one SM (sm_100a).
// Question this answers for kernel (b): what does a "broadcast" LDS.128 / LDS.64 cost when the lanes of a warp
// read 1, 2, 4, 8 ... distinct chunks, and does it matter whether the distinct chunks alternate by lane or are
// split by half / quarter warp?  Output: SM cycles per warp-level load instruction with 16 warps in flight
// (the pipe is saturated, so this is the reciprocal throughput = wavefronts per instruction).
//
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lds_wavefronts lds_wavefronts.cu && ./lds_wavefronts
#include <cstdio>
#include <cuda_runtime.h>

template <int W>
__device__ __forceinline__ void lds(unsigned addr, float (&acc)[4], bool use) {
    float a, b, c, d;
    if constexpr (W == 4) {
        asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(a) : "r"(addr) : "memory");
        if (use) acc[0] += a;
    } else if constexpr (W == 8) {
        asm volatile("ld.volatile.shared.v2.f32 {%0,%1}, [%2];" : "=f"(a), "=f"(b) : "r"(addr) : "memory");
        if (use) { acc[0] += a; acc[1] += b; }
    } else {
        asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(a), "=f"(b), "=f"(c), "=f"(d) : "r"(addr) : "memory");
        if (use) { acc[0] += a; acc[1] += b; acc[2] += c; acc[3] += d; }
    }
}

// chunk index of a lane: pattern selects how many distinct chunks (nd) and how they map to lanes
//   mode 0: chunk = lane % nd  (alternating)      mode 1: chunk = lane / (32/nd)  (blocked)
struct LaneOffsets { int off[32]; };

template <int W>
__global__ void probe(LaneOffsets lo, int iters, long long* cycles, float* sink) {
    extern __shared__ __align__(16) float sm[];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = 1.f;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    unsigned base = (unsigned)__cvta_generic_to_shared(sm) + lo.off[lane];
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 16; ++u) lds<W>(base + u * 1024, acc, u == 15);   // 16 independent volatile loads, 1 KiB apart (same banks); only the last is consumed
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) *cycles = t1 - t0;
    sink[threadIdx.x] = acc[0] + acc[1] + acc[2] + acc[3];
}

template <int W>
static double measure(const LaneOffsets& lo, long long* d_cyc, float* d_sink) {
    const int warps = 16, iters = 2000;
    probe<W><<<1, 32 * warps, 32768>>>(lo, 10, d_cyc, d_sink);   // warm-up
    probe<W><<<1, 32 * warps, 32768>>>(lo, iters, d_cyc, d_sink);
    long long c = 0;
    cudaMemcpy(&c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost);
    return (double)c / ((double)iters * 16 * warps);
}

template <int W>
static void run(const char* name, int nd, int mode, int stride, long long* d_cyc, float* d_sink) {
    LaneOffsets lo;
    for (int lane = 0; lane < 32; ++lane) lo.off[lane] = (mode == 0 ? lane % nd : lane / (32 / nd)) * stride;
    printf("%-8s distinct=%2d %-11s stride=%4dB  %.2f cycles/instr\n", name, nd, mode == 0 ? "alternating" : "blocked",
           stride, measure<W>(lo, d_cyc, d_sink));
}

// arbitrary pattern: offset(lane) = (lane / div % mod) * stride  [+ (lane / div2 % mod2) * stride2]
template <int W>
static void run2(const char* what, int div, int mod, int stride, int div2, int mod2, int stride2, long long* d_cyc,
                 float* d_sink) {
    LaneOffsets lo;
    for (int lane = 0; lane < 32; ++lane)
        lo.off[lane] = (lane / div % mod) * stride + (mod2 ? (lane / div2 % mod2) * stride2 : 0);
    printf("LDS.%-3d %-58s %.2f cycles/instr\n", W * 8, what, measure<W>(lo, d_cyc, d_sink));
}

int main() {
    long long* d_cyc; float* d_sink;
    cudaMalloc(&d_cyc, sizeof(long long));
    cudaMalloc(&d_sink, 1024 * sizeof(float));
    const int nds[] = {1, 2, 4, 8, 16, 32};
    for (int nd : nds) for (int mode = 0; mode < 2; ++mode) {
        if (nd == 1 && mode == 1) continue;
        run<16>("LDS.128", nd, mode, 16, d_cyc, d_sink);
    }
    run<16>("LDS.128", 2, 0, 64, d_cyc, d_sink);      // two chunks in different bank groups
    run<16>("LDS.128", 2, 0, 128, d_cyc, d_sink);     // two chunks on the same banks (conflict)
    run<16>("LDS.128", 4, 0, 32, d_cyc, d_sink);
    for (int nd : nds) for (int mode = 0; mode < 2; ++mode) {
        if (nd == 1 && mode == 1) continue;
        run<8>("LDS.64", nd, mode, 8, d_cyc, d_sink);
    }
    run<8>("LDS.64", 4, 0, 32, d_cyc, d_sink);        // kernel (b) quad mapping: float2 pieces 8 floats apart
    for (int nd : nds) for (int mode = 0; mode < 2; ++mode) {
        if (nd == 1 && mode == 1) continue;
        run<4>("LDS.32", nd, mode, 4, d_cyc, d_sink);
    }
    // patterns of kernel (b) (LP = 20: CT = 5 latent float4 groups, 6 ETF groups; HS = 24 floats, XS = 28 floats)
    run2<16>("grads dWd A: H chunk lt = lane/6 (5 groups of 6 lanes)", 6, 6, 16, 1, 0, 0, d_cyc, d_sink);
    run2<16>("grads dWd B: dZy chunk et = lane%6", 1, 6, 16, 1, 0, 0, d_cyc, d_sink);
    run2<16>("grads dWe A: X chunk et = lane/5", 5, 8, 16, 1, 0, 0, d_cyc, d_sink);
    run2<16>("grads dWe B: dZh chunk lt = lane%5", 1, 5, 16, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows x: 16 rows (pairs share), row stride 112 B", 2, 16, 112, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows x: 16 rows (pairs share), row stride 96 B", 2, 16, 96, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows x: 16 rows (pairs share), row stride 80 B", 2, 16, 80, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows x: 16 rows, lanes l and l+16 share, stride 112 B", 1, 16, 112, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows x: 8 rows (quads share), stride 112 B", 4, 8, 112, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows x: 8 rows, lane%8, stride 112 B", 1, 8, 112, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows x: 32 rows, stride 112 B", 1, 32, 112, 1, 0, 0, d_cyc, d_sink);
    run2<16>("rows own cols: 16 rows x 2 parts (row stride 96 B + part*16)", 2, 16, 96, 1, 2, 16, d_cyc, d_sink);
    run2<16>("4 groups of 8 lanes x 8 alternating chunks", 8, 4, 128, 1, 8, 16, d_cyc, d_sink);
    run2<16>("16 pairs, arbitrary: pair p -> chunk 3p", 2, 16, 48, 1, 0, 0, d_cyc, d_sink);
    run2<16>("16 pairs, arbitrary: pair p -> chunk 5p", 2, 16, 80, 1, 0, 0, d_cyc, d_sink);
    run2<16>("8 quads -> chunk q", 4, 8, 16, 1, 0, 0, d_cyc, d_sink);
    run2<16>("lanes 0-15 chunk l, lanes 16-31 same", 1, 16, 16, 1, 0, 0, d_cyc, d_sink);
    run2<8>("quad weights: part = lane%4 (8 B apart)", 1, 4, 8, 1, 0, 0, d_cyc, d_sink);
    run2<8>("quad weights: part = lane/8 (8 B apart)", 8, 4, 8, 1, 0, 0, d_cyc, d_sink);
    run2<8>("16 pairs -> 8 B chunk p", 2, 16, 8, 1, 0, 0, d_cyc, d_sink);
    run2<8>("lanes l, l+16 share 8 B chunk l", 1, 16, 8, 1, 0, 0, d_cyc, d_sink);
    run2<8>("8 rows (lane%8) stride 112 B + part (lane/8) * 8", 1, 8, 112, 8, 4, 8, d_cyc, d_sink);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
