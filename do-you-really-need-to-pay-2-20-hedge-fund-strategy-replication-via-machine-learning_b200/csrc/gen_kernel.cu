// Generator inference: the saved LSTM generator of the reference (GAN/MTSS_WGAN_GP.py:221-235 `build_generator`; the
// model file the notebook loads at autoencoder_v4.ipynb:1277 and runs at :1291) as one fused forward kernel —
//   LSTM(units, activation = recurrent_activation = sigmoid, return_sequences) [-> LeakyReLU] -> LayerNormalization ->
//   LSTM(...) [-> LeakyReLU] -> LayerNormalization -> Dense(F_out)
// — producing the (n, T, F_out) paths that become configs[4]'s augmented training tasks without a host round trip.
// Keras 2.7 semantics restated (parity unpinned: Keras is not installable; oracle/gen_oracle.py is the numpy restatement):
//   z = x_t W + h_{t-1} U + b, gates in Keras' order i, f, c, o;  c_t = f c_{t-1} + i act(z_c);  h_t = o act(c_t)
//   LayerNormalization over the last axis: (x - mean) * rsqrt(var + eps) * gamma + beta, biased variance.
// One CTA of 4*U threads (one per gate column) walks P paths (4, 8, 16 or 20) through all T steps; the weights (552 KB for the
// reference's 36 -> 100 -> 100 -> 36 model) stream from L2 once per step and CTA, every load feeding P multiply-adds.
#include "common.cuh"

#include <cstdlib>

namespace hfrep {

struct GenParams {
    const float* W;       // packed weights (see hfrep.h)
    const float* noise;   // [n][T][F_in]
    float* out;           // [n][T][F_out]
    int n, T, F_in, U, F_out, lrelu1, lrelu2;
    float alpha, ln_eps;
};

__device__ __forceinline__ float sigmoidf_full(float x) { return 1.f / (1.f + expf(-x)); }

// Shared-memory layout: activations are stored [feature][PS] with the CTA's P paths fastest (PS = P + 4 floats: a row
// stride of 4 mod 8 words, so threads writing consecutive rows with 16-byte stores hit different bank groups).  A gate
// column's dot product then reads FOUR paths of one input feature with one broadcast LDS.128 and feeds two packed FFMA2
// with the weight it loaded once: per four features 4 LDG + P LDS.128 + 2P FFMA2 for 4P multiply-adds per thread (the first
// version read one path per LDS: 4 + 4P + 4P).  Summation order per output is unchanged (bias, inputs, recurrent inputs,
// ascending).
template <int P>
__global__ void __launch_bounds__(512, 2) generator_kernel(const GenParams p) {
    extern __shared__ __align__(16) float sm[];
    constexpr int PS = P + 4, P4 = P / 4;
    const int U = p.U, G = 4 * U, F_in = p.F_in, F_out = p.F_out, T = p.T;
    const int tid = threadIdx.x, nth = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = nth >> 5;
    const float alpha = p.alpha;
    float* x = sm;                    // [F_in][PS]  this step's noise rows
    float* h1 = x + F_in * PS;        // [U][PS]
    float* a1 = h1 + U * PS;          // [U][PS]     LayerNorm(LeakyReLU(h1))
    float* h2 = a1 + U * PS;          // [U][PS]
    float* a2 = h2 + U * PS;          // [U][PS]
    float* c1 = a2 + U * PS;          // [U][PS]     cell states (shared, not registers: two CTAs of 64-register threads per SM)
    float* c2 = c1 + U * PS;          // [U][PS]
    float* z = c2 + U * PS;           // [P][G]      gate pre-activations
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

    for (int i = tid; i < (F_in + 6 * U) * PS; i += nth) sm[i] = 0.f;
    __syncthreads();

    // acc[q] += sum_k in[k][q] * Wm[k][col]: four features per round (the four weight loads in flight together)
    auto accumulate = [&](float2 (&acc)[P / 2], const float* in, int n_in, const float* Wm) {
        const float* wp = Wm + tid;
        int k = 0;
        for (; k + 4 <= n_in; k += 4) {
            float w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) w[i] = __ldg(wp + (size_t)(k + i) * G);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 ww = make_float2(w[i], w[i]);
#pragma unroll
                for (int q4 = 0; q4 < P4; ++q4) {
                    const float4 v = *reinterpret_cast<const float4*>(in + (k + i) * PS + 4 * q4);
                    acc[2 * q4] = __ffma2_rn(make_float2(v.x, v.y), ww, acc[2 * q4]);
                    acc[2 * q4 + 1] = __ffma2_rn(make_float2(v.z, v.w), ww, acc[2 * q4 + 1]);
                }
            }
        }
        for (; k < n_in; ++k) {
            const float w = __ldg(wp + (size_t)k * G);
            const float2 ww = make_float2(w, w);
#pragma unroll
            for (int q4 = 0; q4 < P4; ++q4) {
                const float4 v = *reinterpret_cast<const float4*>(in + k * PS + 4 * q4);
                acc[2 * q4] = __ffma2_rn(make_float2(v.x, v.y), ww, acc[2 * q4]);
                acc[2 * q4 + 1] = __ffma2_rn(make_float2(v.z, v.w), ww, acc[2 * q4 + 1]);
            }
        }
    };
    // z[q][col] = bias[col] + sum_k in[k][q] Win[k][col] + sum_k hprev[k][q] Urec[k][col]   (col = tid)
    auto gates = [&](const float* in, int n_in, const float* Win, const float* hprev, const float* Urec, const float* bias) {
        if (tid < G) {
            float2 acc[P / 2];
            const float b = __ldg(bias + tid);
#pragma unroll
            for (int q2 = 0; q2 < P / 2; ++q2) acc[q2] = make_float2(b, b);
            accumulate(acc, in, n_in, Win);
            accumulate(acc, hprev, U, Urec);
#pragma unroll
            for (int q2 = 0; q2 < P / 2; ++q2) { z[(2 * q2) * G + tid] = acc[q2].x; z[(2 * q2 + 1) * G + tid] = acc[q2].y; }
        }
    };
    // gate combination for unit tid, then [LeakyReLU +] LayerNormalization of each path's U outputs (a warp per four paths)
    auto cell_and_norm = [&](float* c, float* h, float* a, int lrelu, const float* gamma, const float* beta) {
        if (tid < U) {
#pragma unroll
            for (int q4 = 0; q4 < P4; ++q4) {
                float hv[4];
                float4* cp = reinterpret_cast<float4*>(c + tid * PS + 4 * q4);
                const float4 c4 = *cp;
                float cv[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int q = 4 * q4 + i;
                    const float* zq = z + q * G;
                    const float ig = sigmoidf_full(zq[tid]), fg = sigmoidf_full(zq[U + tid]);
                    const float cand = sigmoidf_full(zq[2 * U + tid]), og = sigmoidf_full(zq[3 * U + tid]);
                    cv[i] = fg * cv[i] + ig * cand;
                    hv[i] = og * sigmoidf_full(cv[i]);
                }
                *cp = make_float4(cv[0], cv[1], cv[2], cv[3]);
                *reinterpret_cast<float4*>(h + tid * PS + 4 * q4) = make_float4(hv[0], hv[1], hv[2], hv[3]);
            }
        }
        __syncthreads();
        auto act = [&](float v) { return (lrelu && !(v > 0.f)) ? alpha * v : v; };
        for (int q4 = warp; q4 < P4; q4 += nwarp) {
            float s[4] = {0.f, 0.f, 0.f, 0.f};
            for (int u = lane; u < U; u += 32) {
                const float4 v = *reinterpret_cast<const float4*>(h + u * PS + 4 * q4);
                s[0] += act(v.x); s[1] += act(v.y); s[2] += act(v.z); s[3] += act(v.w);
            }
            float mean[4], inv[4], ss[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 4; ++i) mean[i] = warp_sum(s[i]) / (float)U;
            for (int u = lane; u < U; u += 32) {
                const float4 v = *reinterpret_cast<const float4*>(h + u * PS + 4 * q4);
                const float d0 = act(v.x) - mean[0], d1 = act(v.y) - mean[1], d2 = act(v.z) - mean[2], d3 = act(v.w) - mean[3];
                ss[0] += d0 * d0; ss[1] += d1 * d1; ss[2] += d2 * d2; ss[3] += d3 * d3;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) inv[i] = rsqrtf(warp_sum(ss[i]) / (float)U + p.ln_eps);
            for (int u = lane; u < U; u += 32) {
                const float4 v = *reinterpret_cast<const float4*>(h + u * PS + 4 * q4);
                const float ga = __ldg(gamma + u), be = __ldg(beta + u);
                *reinterpret_cast<float4*>(a + u * PS + 4 * q4) =
                    make_float4((act(v.x) - mean[0]) * inv[0] * ga + be, (act(v.y) - mean[1]) * inv[1] * ga + be,
                                (act(v.z) - mean[2]) * inv[2] * ga + be, (act(v.w) - mean[3]) * inv[3] * ga + be);
            }
        }
        __syncthreads();
    };

    for (int t = 0; t < T; ++t) {
        for (int i = tid; i < P * F_in; i += nth) {
            const int q = i / F_in, k = i - q * F_in;
            x[k * PS + q] = q < np ? __ldg(p.noise + ((size_t)(path0 + q) * T + t) * F_in + k) : 0.f;
        }
        __syncthreads();
        gates(x, F_in, W1, h1, U1, b1);
        __syncthreads();
        cell_and_norm(c1, h1, a1, p.lrelu1, g1, be1);
        gates(a1, U, W2, h2, U2, b2);
        __syncthreads();
        cell_and_norm(c2, h2, a2, p.lrelu2, g2, be2);
        for (int i = tid; i < P4 * F_out; i += nth) {            // Dense: out = a2 Kd + bd, four paths per thread
            const int q4 = i / F_out, j = i - q4 * F_out;
            const float b = __ldg(bd + j);
            float2 acc0 = make_float2(b, b), acc1 = acc0;
            for (int u = 0; u < U; ++u) {
                const float w = __ldg(Kd + (size_t)u * F_out + j);
                const float4 v = *reinterpret_cast<const float4*>(a2 + u * PS + 4 * q4);
                acc0 = __ffma2_rn(make_float2(v.x, v.y), make_float2(w, w), acc0);
                acc1 = __ffma2_rn(make_float2(v.z, v.w), make_float2(w, w), acc1);
            }
            const float o[4] = {acc0.x, acc0.y, acc1.x, acc1.y};
#pragma unroll
            for (int r = 0; r < 4; ++r)
                if (4 * q4 + r < np) p.out[((size_t)(path0 + 4 * q4 + r) * T + t) * F_out + j] = o[r];
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
    const size_t smem = sizeof(float) * ((size_t)(p.F_in + 6 * p.U) * (P + 4) + (size_t)P * 4 * p.U);
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
    int P = n_paths >= 40 * sms ? 20 : (n_paths >= 16 * sms ? 16 : (n_paths >= 4 * sms ? 8 : 4));   // (100 000 paths: 20 -> 408 k paths/s,
                                                                                              //  16 -> 384 k, 24 -> 376 k, 32 (one CTA per SM) -> 322 k)
    if (const char* e = getenv("HFREP_GEN_P")) P = atoi(e);                                  // (tuning only)
    if (P == 20) return launch_generator<20>(p, (cudaStream_t)stream);
    if (P == 16) return launch_generator<16>(p, (cudaStream_t)stream);
    if (P == 8) return launch_generator<8>(p, (cudaStream_t)stream);
    return launch_generator<4>(p, (cudaStream_t)stream);
}
