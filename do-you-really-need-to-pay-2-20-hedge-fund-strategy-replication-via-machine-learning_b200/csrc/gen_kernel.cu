// Generator inference: the saved LSTM generator of the reference (GAN/MTSS_WGAN_GP.py:221-235 `build_generator`; the
// model file the notebook loads at autoencoder_v4.ipynb:1277 and runs at :1291) as one fused forward kernel —
//   LSTM(units, activation = recurrent_activation = sigmoid, return_sequences) [-> LeakyReLU] -> LayerNormalization ->
//   LSTM(...) [-> LeakyReLU] -> LayerNormalization -> Dense(F_out)
// — producing the (n, T, F_out) paths that become configs[4]'s augmented training tasks without a host round trip.
// Keras 2.7 semantics restated (parity unpinned: Keras is not installable; oracle/gen_oracle.py is the numpy restatement):
//   z = x_t W + h_{t-1} U + b, gates in Keras' order i, f, c, o;  c_t = f c_{t-1} + i act(z_c);  h_t = o act(c_t)
//   LayerNormalization over the last axis: (x - mean) * rsqrt(var + eps) * gamma + beta, biased variance.
// One CTA of 4*U threads (one per gate column) walks P paths through all T steps; the weights (552 KB for the reference's
// 36 -> 100 -> 100 -> 36 model) stream from L2 once per step and CTA, every load feeding P multiply-adds.
#include "common.cuh"

namespace hfrep {

struct GenParams {
    const float* W;       // packed weights (see hfrep.h)
    const float* noise;   // [n][T][F_in]
    float* out;           // [n][T][F_out]
    int n, T, F_in, U, F_out, lrelu1, lrelu2;
    float alpha, ln_eps;
};

__device__ __forceinline__ float sigmoidf_full(float x) { return 1.f / (1.f + expf(-x)); }

template <int P>
__global__ void __launch_bounds__(512) generator_kernel(const GenParams p) {
    extern __shared__ __align__(16) float sm[];
    const int U = p.U, G = 4 * U, F_in = p.F_in, F_out = p.F_out, T = p.T;
    const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = nth >> 5;
    const float alpha = p.alpha;
    float* x = sm;                    // [P][F_in]   this step's noise rows
    float* h1 = x + P * F_in;         // [P][U]
    float* a1 = h1 + P * U;           // [P][U]      LayerNorm(LeakyReLU(h1))
    float* h2 = a1 + P * U;           // [P][U]
    float* a2 = h2 + P * U;           // [P][U]
    float* z = a2 + P * U;            // [P][G]      gate pre-activations
    const float* W1 = p.W;                          // [F_in][G]
    const float* U1 = W1 + (size_t)F_in * G;        // [U][G]
    const float* b1 = U1 + (size_t)U * G;           // [G]
    const float* g1 = b1 + G;                       // [U]
    const float* be1 = g1 + U;                      // [U]
    const float* W2 = be1 + U;                      // [U][G]
    const float* U2 = W2 + (size_t)U * G;           // [U][G]
    const float* b2 = U2 + (size_t)U * G;           // [G]
    const float* g2 = b2 + G;
    const float* be2 = g2 + U;
    const float* Kd = be2 + U;                      // [U][F_out]
    const float* bd = Kd + (size_t)U * F_out;       // [F_out]
    const int path0 = blockIdx.x * P;
    const int np = min(P, p.n - path0);

    for (int i = tid; i < P * U; i += nth) h1[i] = h2[i] = 0.f;
    float c1[P], c2[P];               // cell states of unit `tid` (threads tid < U)
#pragma unroll
    for (int q = 0; q < P; ++q) c1[q] = c2[q] = 0.f;
    __syncthreads();

    // z[q][col] = bias[col] + sum_k in[q][k] Win[k][col] + sum_k hprev[q][k] Urec[k][col]   (col = tid)
    auto gates = [&](const float* in, int n_in, const float* Win, const float* hprev, const float* Urec, const float* bias) {
        if (tid < G) {
            float acc[P];
            const float b = __ldg(bias + tid);
#pragma unroll
            for (int q = 0; q < P; ++q) acc[q] = b;
            for (int k = 0; k < n_in; ++k) {
                const float w = __ldg(Win + (size_t)k * G + tid);
#pragma unroll
                for (int q = 0; q < P; ++q) acc[q] = fmaf(in[q * n_in + k], w, acc[q]);
            }
            for (int k = 0; k < U; ++k) {
                const float w = __ldg(Urec + (size_t)k * G + tid);
#pragma unroll
                for (int q = 0; q < P; ++q) acc[q] = fmaf(hprev[q * U + k], w, acc[q]);
            }
#pragma unroll
            for (int q = 0; q < P; ++q) z[q * G + tid] = acc[q];
        }
    };
    // gate combination for unit tid, then [LeakyReLU +] LayerNormalization of each path's U outputs (a warp per path)
    auto cell_and_norm = [&](float (&c)[P], float* h, float* a, int lrelu, const float* gamma, const float* beta) {
        if (tid < U) {
#pragma unroll
            for (int q = 0; q < P; ++q) {
                const float* zq = z + q * G;
                const float ig = sigmoidf_full(zq[tid]), fg = sigmoidf_full(zq[U + tid]);
                const float cand = sigmoidf_full(zq[2 * U + tid]), og = sigmoidf_full(zq[3 * U + tid]);
                c[q] = fg * c[q] + ig * cand;
                h[q * U + tid] = og * sigmoidf_full(c[q]);
            }
        }
        __syncthreads();
        auto act = [&](float v) { return (lrelu && !(v > 0.f)) ? alpha * v : v; };
        for (int q = warp; q < P; q += nwarp) {
            float s = 0.f;
            for (int u = lane; u < U; u += 32) s += act(h[q * U + u]);
            const float mean = warp_sum(s) / (float)U;
            float ss = 0.f;
            for (int u = lane; u < U; u += 32) { const float d = act(h[q * U + u]) - mean; ss += d * d; }
            const float inv = rsqrtf(warp_sum(ss) / (float)U + p.ln_eps);
            for (int u = lane; u < U; u += 32)
                a[q * U + u] = (act(h[q * U + u]) - mean) * inv * __ldg(gamma + u) + __ldg(beta + u);
        }
        __syncthreads();
    };

    for (int t = 0; t < T; ++t) {
        for (int i = tid; i < P * F_in; i += nth) {
            const int q = i / F_in, k = i - q * F_in;
            x[i] = q < np ? __ldg(p.noise + ((size_t)(path0 + q) * T + t) * F_in + k) : 0.f;
        }
        __syncthreads();
        gates(x, F_in, W1, h1, U1, b1);
        __syncthreads();
        cell_and_norm(c1, h1, a1, p.lrelu1, g1, be1);
        gates(a1, U, W2, h2, U2, b2);
        __syncthreads();
        cell_and_norm(c2, h2, a2, p.lrelu2, g2, be2);
        for (int i = tid; i < P * F_out; i += nth) {             // Dense: out = a2 Kd + bd
            const int q = i / F_out, j = i - q * F_out;
            float acc = __ldg(bd + j);
            for (int u = 0; u < U; ++u) acc = fmaf(a2[q * U + u], __ldg(Kd + (size_t)u * F_out + j), acc);
            if (q < np) p.out[((size_t)(path0 + q) * T + t) * F_out + j] = acc;
        }
        __syncthreads();
    }
}

}  // namespace hfrep

using namespace hfrep;

extern "C" size_t hfrep_generator_weight_count(int F_in, int units, int F_out) {
    if (F_in <= 0 || units <= 0 || F_out <= 0) return 0;
    const size_t G = 4 * (size_t)units;
    return (size_t)F_in * G + units * G + G + 2 * units + 2 * (units * G) + G + 2 * units + (size_t)units * F_out + F_out;
}

template <int P>
static int launch_generator(const GenParams& p, cudaStream_t stream) {
    const int threads = ((4 * p.U + 31) / 32) * 32;
    const size_t smem = sizeof(float) * ((size_t)P * p.F_in + 4 * (size_t)P * p.U + (size_t)P * 4 * p.U);
    HFREP_REQUIRE(smem <= 200 * 1024, "generator_forward: %zu bytes of shared memory needed", smem);
    HFREP_CUDA_OK(allow_max_dynamic_smem(generator_kernel<P>));
    generator_kernel<P><<<(p.n + P - 1) / P, threads, smem, stream>>>(p);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_generator_forward(const float* weights, int F_in, int units, int F_out, int lrelu_after_lstm1,
                                       int lrelu_after_lstm2, float alpha, float ln_eps, const float* noise, int n_paths,
                                       int T, float* out, void* stream) {
    HFREP_REQUIRE(weights && noise && out, "generator_forward: null argument");
    HFREP_REQUIRE(n_paths >= 0 && T >= 1, "generator_forward: n_paths=%d T=%d", n_paths, T);
    if (n_paths == 0) return HFREP_OK;
    HFREP_REQUIRE(units >= 1 && units <= 128 && F_in >= 1 && F_in <= 128 && F_out >= 1 && F_out <= 128,
                  "generator_forward: F_in=%d units=%d F_out=%d outside 1..128", F_in, units, F_out);
    GenParams p;
    p.W = weights; p.noise = noise; p.out = out; p.n = n_paths; p.T = T; p.F_in = F_in; p.U = units; p.F_out = F_out;
    p.lrelu1 = lrelu_after_lstm1; p.lrelu2 = lrelu_after_lstm2; p.alpha = alpha; p.ln_eps = ln_eps;
    // paths per CTA: enough to amortise the weight stream once there are many paths, few when there are few (more CTAs)
    const int sms = sm_count();
    if (n_paths >= 16 * sms) return launch_generator<16>(p, (cudaStream_t)stream);
    if (n_paths >= 4 * sms) return launch_generator<8>(p, (cudaStream_t)stream);
    return launch_generator<2>(p, (cudaStream_t)stream);
}
