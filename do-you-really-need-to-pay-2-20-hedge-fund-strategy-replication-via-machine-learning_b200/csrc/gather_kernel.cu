// Kernel (a): window gather (+ optional per-window MinMax scaling), rolling ex-post volatility,
// and the FP32-FFMA peak micro-benchmark that provides kernel (b)'s roofline denominator.
//
// window_gather replaces the python slicing loops of helper.random_sampling (helper.py:55-58),
// helper.factor_hf_split (helper.py:145-152) and the rolling slices of Autoencoder_encapsulate.py:149-150;
// with minmax=1 it also applies MinMaxScaler().fit_transform (:62-65) per window.
// HBM-bound: n*len*ncols*4 bytes written, the same read (re-reads of overlapping windows hit L2).
#include "common.cuh"

namespace hfrep {

// One CTA per window per pass (grid-stride over windows).  Threads walk the (row, col) elements of the
// window in output order, so stores are fully coalesced and loads are coalesced per source row.
// Fast path: ncols % 4 == 0 and 16-byte aligned rows -> float4 loads/stores.
// TS = float or double: the returns matrix as the caller holds it.  With a float64 source (the reference's DataFrames are
// float64) min, max and the scaling are taken on the float64 values and the result is rounded to float32 once — exactly
// MinMaxScaler on float64 followed by Keras' cast (Autoencoder_encapsulate.py:62-65, :83).
template <typename TS>
__global__ void __launch_bounds__(256) window_gather_kernel(const TS* __restrict__ R, int C,
                                                            const int* __restrict__ starts, int n, int len,
                                                            int col_lo, int ncols, int minmax, int vec4,
                                                            float* __restrict__ out) {
    extern __shared__ double cst[];          // [2][ncols]: scale, shift (only when minmax)
    for (int w = blockIdx.x; w < n; w += gridDim.x) {
        const TS* src = R + (size_t)starts[w] * C + col_lo;
        float* dst = out + (size_t)w * len * ncols;
        if (minmax) {
            __syncthreads();
            for (int c = threadIdx.x; c < ncols; c += blockDim.x) {
                TS lo = INFINITY, hi = -INFINITY;
                for (int r = 0; r < len; ++r) {
                    const TS v = __ldg(src + (size_t)r * C + c);
                    lo = v < lo ? v : lo; hi = v > hi ? v : hi;
                }
                double rng = (double)hi - (double)lo;
                if (rng == 0.) rng = 1.;
                const double scale = 1. / rng;
                cst[c] = scale;
                cst[ncols + c] = __dsub_rn(0., __dmul_rn((double)lo, scale));
            }
            __syncthreads();
            for (int idx = threadIdx.x; idx < len * ncols; idx += blockDim.x) {
                const int r = idx / ncols, c = idx - r * ncols;
                dst[idx] = (float)__dadd_rn(__dmul_rn((double)__ldg(src + (size_t)r * C + c), cst[c]), cst[ncols + c]);   // X *= scale; X += min_ (no FMA contraction)
            }
        } else if (sizeof(TS) == 4 && vec4) {    // (float source only)
            const int q = ncols >> 2;
            for (int idx = threadIdx.x; idx < len * q; idx += blockDim.x) {
                const int r = idx / q, c4 = idx - r * q;
                const float4 v = __ldg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(src) + (size_t)r * C) + c4);
                reinterpret_cast<float4*>(dst)[idx] = v;
            }
        } else {
            for (int idx = threadIdx.x; idx < len * ncols; idx += blockDim.x) {
                const int r = idx / ncols, c = idx - r * ncols;
                dst[idx] = (float)__ldg(src + (size_t)r * C + c);
            }
        }
    }
}

// sigma[i][e] = sqrt(var(R[start_i : start_i+window, e], ddof=1)) — the diagonal of DataFrame.cov() (helper.py:121)
__global__ void rolling_sigma_kernel(const double* __restrict__ R, int E, const int* __restrict__ starts, int n,
                                     int window, double* __restrict__ sigma) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * E) return;
    const int i = idx / E, e = idx - i * E;
    const double* src = R + (size_t)starts[i] * E + e;
    double mean = 0.;
    for (int r = 0; r < window; ++r) mean += src[(size_t)r * E];
    mean /= window;
    double ss = 0.;
    for (int r = 0; r < window; ++r) { const double d = src[(size_t)r * E] - mean; ss += d * d; }
    sigma[idx] = sqrt(ss / (window - 1));
}

// helper.ex_post_return for arbitrary weights: one thread per (month, fund)
__global__ void ex_post_kernel(const double* __restrict__ ante, const double* __restrict__ w,
                               const double* __restrict__ sigma, int T, int F, int E, double param, double phi,
                               double* __restrict__ post) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= T * F) return;
    const int t = idx / F, f = idx - t * F;
    double pen = 0.;
    if (t >= 1) {
        const double* wn = w + ((size_t)f * T + t) * E;
        const double* wo = wn - E;
        for (int e = 0; e < E; ++e) {
            const double sg = sigma[(size_t)(t - 1) * E + e] * param;
            const double d = wo[e] - wn[e];
            const double tc = .5 * d * d * sg;                                           // helper.py:80
            const double pi = phi * wn[e] * sg * d - wo[e] * sg * d - .5 * d * d * sg;   // helper.py:92
            pen += tc + pi;
        }
    }
    post[idx] = ante[idx] + pen;
}

// FP32 peak: 8 independent packed-FMA chains per thread (FFMA2 = fma.rn.f32x2, two IEEE FP32 FMAs per issue slot,
// the instruction kernel (b) is built from), `iters` rounds: 32*iters flop per thread.
__global__ void __launch_bounds__(256) ffma_peak_kernel(int iters, float* sink) {
    float2 a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = make_float2(threadIdx.x * 1e-3f + j, 0.5f * j + 1.f);
    const float2 m = make_float2(0.999f, 0.998f), c = make_float2(1e-3f, 2e-3f);
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = __ffma2_rn(a[j], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j].x + a[j].y;
    if (s == 123456.789f) sink[0] = s;        // never true; keeps the chains alive
}

}  // namespace hfrep

using namespace hfrep;

namespace hfrep {
template <typename TS>
static int launch_window_gather_t(const TS* R, int T, int C, const int32_t* starts, int n, int len, int col_lo, int ncols,
                                  int minmax, float* out, cudaStream_t stream) {
    HFREP_REQUIRE(n >= 0 && len >= 0, "window_gather: n=%d len=%d", n, len);
    if (n == 0 || len == 0) return HFREP_OK;             // empty batch: nothing to read or write
    HFREP_REQUIRE(R && starts && out, "window_gather: null argument");
    HFREP_REQUIRE(col_lo >= 0 && ncols >= 1 && col_lo + ncols <= C, "window_gather: columns [%d,%d) outside 0..%d", col_lo,
                  col_lo + ncols, C);
    HFREP_REQUIRE(len <= T, "window_gather: len=%d exceeds T=%d", len, T);
    const int vec4 = sizeof(TS) == 4 && !minmax && (ncols % 4 == 0) && (C % 4 == 0) && (col_lo % 4 == 0) &&
                     (((uintptr_t)R | (uintptr_t)out) & 15) == 0;
    const size_t smem = minmax ? sizeof(double) * 2 * ncols : 0;
    HFREP_REQUIRE(smem <= 48 * 1024, "window_gather: ncols=%d too wide for the minmax path", ncols);
    const int sms = sm_count();
    const int grid = n < 8 * sms ? n : 8 * sms;      // whole multiples of the SM count once there is enough work
    window_gather_kernel<TS><<<grid, 256, smem, stream>>>(R, C, starts, n, len, col_lo, ncols, minmax, vec4, out);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

int launch_window_gather(const float* R, int T, int C, const int32_t* starts, int n, int len, int col_lo, int ncols,
                         int minmax, float* out, cudaStream_t stream) {
    return launch_window_gather_t<float>(R, T, C, starts, n, len, col_lo, ncols, minmax, out, stream);
}
}  // namespace hfrep

extern "C" int hfrep_window_gather_f64(const double* R, int T, int C, const int32_t* starts, int n, int len, int col_lo,
                                       int ncols, int minmax, float* out, void* stream) {
    return launch_window_gather_t<double>(R, T, C, starts, n, len, col_lo, ncols, minmax, out, (cudaStream_t)stream);
}

extern "C" int hfrep_window_gather(const float* R, int T, int C, const int32_t* starts, int n, int len, int col_lo,
                                   int ncols, int minmax, float* out, void* stream) {
    return launch_window_gather(R, T, C, starts, n, len, col_lo, ncols, minmax, out, (cudaStream_t)stream);
}

extern "C" int hfrep_rolling_sigma(const double* R, int T, int E, const int32_t* starts, int n, int window,
                                   double* sigma, void* stream) {
    HFREP_REQUIRE(R && starts && sigma, "rolling_sigma: null argument");
    HFREP_REQUIRE(window >= 2 && window <= T && E >= 1 && n >= 0, "rolling_sigma: window=%d T=%d E=%d n=%d", window, T, E, n);
    if (n == 0) return HFREP_OK;
    const int total = n * E;
    rolling_sigma_kernel<<<(total + 127) / 128, 128, 0, (cudaStream_t)stream>>>(R, E, starts, n, window, sigma);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_ffma_peak(int iters, float* sink, double* flops_host, void* stream) {
    HFREP_REQUIRE(sink && iters > 0, "ffma_peak: bad argument");
    const int blocks = sm_count() * 8, threads = 256;
    ffma_peak_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, sink);
    HFREP_CUDA_OK(cudaGetLastError());
    if (flops_host) *flops_host = 32.0 * iters * (double)blocks * threads;
    return HFREP_OK;
}

extern "C" int hfrep_ex_post_return(const double* ante, const double* weights, const double* sigma, int T, int F, int E,
                                    double param, double phi, double* post, void* stream) {
    HFREP_REQUIRE(ante && weights && sigma && post, "ex_post_return: null argument");
    HFREP_REQUIRE(T >= 1 && F >= 1 && E >= 1, "ex_post_return: T=%d F=%d E=%d", T, F, E);
    const int total = T * F;
    ex_post_kernel<<<(total + 127) / 128, 128, 0, (cudaStream_t)stream>>>(ante, weights, sigma, T, F, E, param, phi, post);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}
