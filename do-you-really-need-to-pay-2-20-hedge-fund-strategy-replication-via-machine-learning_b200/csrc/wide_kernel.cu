// Wide variant of kernels (b) and (c) for the synthetic scale-up (BASELINE configs[3]: 500 ETFs, 10 000 funds,
// 252-day training windows, rolling refit).  Same arithmetic as fit_kernel / eval_kernel — bias-free E -> L -> E
// LeakyReLU autoencoder, MSE, Nadam, validation tail, early stopping (Autoencoder_encapsulate.py:19-35, :72-96), then
// OLS of the funds on the latent factors and the replica return of the next row (:133-201 with the reference's own
// alignment) — for shapes whose state does not fit shared memory: weights and Nadam state live in a caller-provided
// global workspace (L2-resident: 6*E*L floats per task), activations of the current mini-batch in shared memory.
//
// Two fit kernels: fit_wide_kernel (v1: correct first, any E <= 4096 and L <= 64, weights in the global workspace) and
// fit_wide2_kernel (E <= 512, L <= 32: ETF columns dealt to threads, weights in registers, packed FFMA2, shuffle
// scatter-reductions over the ETFs; 4.3-4.6x faster than v1) and, for L <= 8 (configs[3]'s width), fit_wide3_kernel: the
// same ownership, the two contractions over the ETFs as row-mapped register tiles over a shared mini-batch tile (1.55x
// faster than v2).  All FP32 FFMA2: no tensor-core version (DESIGN.md section 4 says why).
#include "common.cuh"

#include <cmath>
#include <cstdlib>

namespace hfrep {

constexpr int WIDE_THREADS = 256;
constexpr int WIDE_MAX_L = 64;

struct WideFitParams {
    const float* X;            // [n_x][T][E] scaled training blocks
    const int* task_x;         // block of each task (NULL: block 0)
    const float* W0;           // [n_tasks][2*E*L]
    const int* perm;           // optional [..][epochs][n_fit]
    long long perm_task_stride;
    const float4* sched;
    float* ws;                 // [n_tasks][6*E*L]: We, Wd, mWe, vWe, mWd, vWd
    float* W_out;              // [n_tasks][2*E*L]
    float* loss_hist; float* val_hist; int* epochs_run; int* status;
    int n_tasks, E, L, T, n_fit, n_val, batch, max_epochs, patience, early_stop;
    float alpha, lr, b1, b2, eps;
};

// One Nadam step of a weight (common.cuh — Keras' operation order, IEEE-exact division and square root: the fast
// sequences, evaluated again with the compiler's range-checked operations if an operand is non-zero and below 2^-64).
__device__ __forceinline__ float nadam_step(float w, float g, float& m, float& v, const float4& sa, float mu1, const float3& rc,
                                            const NadamK& k) {
    nadam_moments(g, m, v, k);
    unsigned lo = 0xffffffffu;
    float d = nadam_decrement<true>(g, m, v, sa, mu1, rc, k, lo);
    if (lo < kNadamSmallBits)
        d = nadam_decrement_checked4(make_float4(g, 0.f, 0.f, 0.f), make_float4(m, 0.f, 0.f, 0.f), make_float4(v, 1.f, 1.f, 1.f), sa, mu1, k).x;
    return __fsub_rn(w, d);
}

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float s = 0.f;
    for (int w = 0; w < WIDE_THREADS / 32; ++w) s += red[w];
    return s;
}

__global__ void __launch_bounds__(WIDE_THREADS) fit_wide_kernel(const WideFitParams p) {
    extern __shared__ __align__(16) float sm[];
    const int task = blockIdx.x, tid = threadIdx.x;
    const int E = p.E, L = p.L, B = p.batch;
    float* H = sm;                         // [B][L]   h = lrelu(x We)
    float* dZh = H + B * L;                // [B][L]
    float* dZy = dZh + B * L;              // [B][E]
    int* rows = reinterpret_cast<int*>(dZy + (size_t)B * E);   // [B] training-block rows of the mini-batch
    float* red = reinterpret_cast<float*>(rows + B);           // [8]

    const size_t EL = (size_t)E * L;
    float* We = p.ws + (size_t)task * 6 * EL;   // [E][L]
    float* Wd = We + EL;                         // [L][E]
    float* mWe = Wd + EL; float* vWe = mWe + EL; float* mWd = vWe + EL; float* vWd = mWd + EL;
    const float* X = p.X + (size_t)(p.task_x ? p.task_x[task] : 0) * p.T * E;
    const float* w0 = p.W0 + (size_t)task * 2 * EL;
    for (size_t i = tid; i < EL; i += WIDE_THREADS) {
        We[i] = w0[i]; Wd[i] = w0[EL + i];
        mWe[i] = vWe[i] = mWd[i] = vWd[i] = 0.f;
    }
    __syncthreads();

    const int* perm_t = p.perm ? p.perm + (size_t)task * p.perm_task_stride : nullptr;
    const float alpha = p.alpha, omb1 = 1.f - p.b1, omb2 = 1.f - p.b2;
    const int n_fit = p.n_fit, n_val = p.n_val;
    int step = 0, wait = 0, epochs_run = 0, st = 0;
    float best = INFINITY;

    // forward of the mini-batch in `rows` (nb rows); train: also dZy, dZh, weight gradients and the Nadam update
    auto run_batch = [&](int nb, bool train) -> float {
        // H = lrelu(X_b We): one output per thread, l fastest (We row coalesced, x broadcast)
        for (int o = tid; o < nb * L; o += WIDE_THREADS) {
            const int r = o / L, l = o - r * L;
            const float* xr = X + (size_t)rows[r] * E;
            float acc = 0.f;
            for (int e = 0; e < E; ++e) acc = fmaf(__ldg(xr + e), We[(size_t)e * L + l], acc);
            H[o] = lrelu(acc, alpha);
        }
        __syncthreads();
        // Zy = H Wd, squared error, dZy: e fastest (Wd row coalesced, h broadcast from shared memory)
        const float coef = 2.f / (float)(nb * E);
        float sse = 0.f;
        for (int o = tid; o < nb * E; o += WIDE_THREADS) {
            const int r = o / E, e = o - r * E;
            float acc = 0.f;
            for (int l = 0; l < L; ++l) acc = fmaf(H[r * L + l], Wd[(size_t)l * E + e], acc);
            const float d = lrelu(acc, alpha) - __ldg(X + (size_t)rows[r] * E + e);
            sse = fmaf(d, d, sse);
            if (train) dZy[o] = lrelu_grad(acc, coef * d, alpha);
        }
        if (!train) return sse;
        __syncthreads();
        // dZh = (dZy Wd^T) o lrelu'(zh): a warp per output, lanes over e (both operands coalesced), fixed-order reduce
        {
            const int lane = tid & 31, warp = tid >> 5;
            for (int o = warp; o < nb * L; o += WIDE_THREADS / 32) {
                const int r = o / L, l = o - r * L;
                float acc = 0.f;
                for (int e = lane; e < E; e += 32) acc = fmaf(dZy[(size_t)r * E + e], Wd[(size_t)l * E + e], acc);
                acc = warp_sum(acc);
                if (lane == 0) dZh[o] = lrelu_grad(H[o], acc, alpha);     // sign(h) == sign(zh) for alpha > 0
            }
        }
        __syncthreads();
        // gradients + Nadam, element-wise over the 2*E*L weights
        const float4 sa = __ldg(p.sched + 2 * step);
        const float mu1 = __ldg(p.sched + 2 * step + 1).x;
        const float3 rc = make_float3(rcp_refined(sa.x), rcp_refined(sa.y), rcp_refined(sa.z));
        const NadamK nk{p.b1, omb1, p.b2, omb2, p.lr, p.eps};
        for (size_t i = tid; i < 2 * EL; i += WIDE_THREADS) {
            float g = 0.f;
            float *w, *m, *v;
            if (i < EL) {                                  // dWd[l][e] = sum_r H[r][l] dZy[r][e]
                const int l = (int)(i / E), e = (int)(i - (size_t)l * E);
                for (int r = 0; r < nb; ++r) g = fmaf(H[r * L + l], dZy[(size_t)r * E + e], g);
                w = Wd + i; m = mWd + i; v = vWd + i;
            } else {                                       // dWe[e][l] = sum_r X_b[r][e] dZh[r][l]
                const size_t j = i - EL;
                const int e = (int)(j / L), l = (int)(j - (size_t)e * L);
                for (int r = 0; r < nb; ++r) g = fmaf(__ldg(X + (size_t)rows[r] * E + e), dZh[r * L + l], g);
                w = We + j; m = mWe + j; v = vWe + j;
            }
            float mm = *m, vv = *v;
            *w = nadam_step(*w, g, mm, vv, sa, mu1, rc, nk);
            *m = mm; *v = vv;
        }
        ++step;
        __syncthreads();
        return sse;
    };

    for (int epoch = 0; epoch < p.max_epochs; ++epoch) {
        float lsum = 0.f, vsum = 0.f;
        for (int lo = 0; lo < n_fit; lo += B) {
            const int nb = min(B, n_fit - lo);
            for (int q = tid; q < nb; q += WIDE_THREADS)
                rows[q] = perm_t ? __ldg(perm_t + (size_t)epoch * n_fit + lo + q) : lo + q;
            __syncthreads();
            lsum += run_batch(nb, true);
        }
        for (int lo = 0; lo < n_val; lo += B) {
            const int nb = min(B, n_val - lo);
            __syncthreads();
            for (int q = tid; q < nb; q += WIDE_THREADS) rows[q] = n_fit + lo + q;
            __syncthreads();
            vsum += run_batch(nb, false);
        }
        const float loss = block_sum(lsum, red) / ((float)E * n_fit);
        const float val = n_val > 0 ? block_sum(vsum, red) / ((float)E * n_val) : NAN;
        if (tid == 0) {
            p.loss_hist[(size_t)task * p.max_epochs + epoch] = loss;
            p.val_hist[(size_t)task * p.max_epochs + epoch] = val;
        }
        epochs_run = epoch + 1;
        if (!isfinite(loss)) st = 1;
        if (p.early_stop && n_val > 0) {         // no validation rows: no val_loss, Keras never stops (see fit_kernel.cu)
            ++wait;
            if (val < best) { best = val; wait = 0; }
            if (wait >= p.patience && epoch > 0) break;
        }
        __syncthreads();
    }
    float* wo = p.W_out + (size_t)task * 2 * EL;
    for (size_t i = tid; i < EL; i += WIDE_THREADS) { wo[i] = We[i]; wo[EL + i] = Wd[i]; }
    if (tid == 0) { p.epochs_run[task] = epochs_run; p.status[task] = st; }
}

// ------------------------------------------------------------------------------------------------ wide fit, v2
// The same training run with the ETF columns dealt to threads: thread t owns columns [t*EPT, t*EPT+EPT) and keeps ITS rows
// of We and ITS columns of Wd in registers for the whole run, so no weight ever moves through shared memory.  Of the five
// contractions of a step, three are then thread-local (zy = h Wd, dWd = h^T dzy, dWe = x^T dzh: the other operand is a
// broadcast row of h / dzh from shared memory) and two reduce over the ETFs, i.e. over threads (zh = x We, dzh = dzy Wd^T):
// a lane forms its 32 partial products (32/LP rows x LP latents), a 31-shuffle scatter-reduction leaves each lane with one
// complete warp sum, the warps' sums meet in shared memory.  x rows are read from global memory (L1/L2) three times per step,
// coalesced; dzy is never stored.  Nadam's m, v live in the caller's workspace, laid out [slot][thread] (coalesced).
// E <= 512, L <= 32 (else the v1 kernel above).
template <int NV, int MASK>
__device__ __forceinline__ float warp_scatter_sum(float (&v)[32], int lane, int& j0) {
    if constexpr (NV == 1) {
        float s = v[0];
#pragma unroll
        for (int m = MASK; m >= 1; m >>= 1) s += __shfl_xor_sync(0xffffffffu, s, m);
        return s;
    } else {
        constexpr int HALF = NV / 2;
        const bool up = (lane & MASK) != 0;
#pragma unroll
        for (int i = 0; i < HALF; ++i) {
            const float send = up ? v[i] : v[i + HALF];
            const float keep = up ? v[i + HALF] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, MASK);
        }
        if (up) j0 += HALF;
        return warp_scatter_sum<HALF, MASK / 2>(v, lane, j0);
    }
}

template <int LP, int EPT, bool MVS>
__global__ void __launch_bounds__(512 / EPT) fit_wide2_kernel(const WideFitParams p) {
    extern __shared__ __align__(16) float sm[];
    constexpr int RG = 32 / LP;                 // batch rows per reduction group (32 values per lane)
    constexpr int LH = LP / 2;                  // latent columns are held in pairs: every FMA is a packed FFMA2
    const int task = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int NT = blockDim.x, NWARP = NT >> 5;
    const int E = p.E, L = p.L, B = p.batch;
    const int BP = (B + 3) & ~3;                // rows of the shared tiles (a multiple of RG)
    float* H = sm;                              // [BP][LP]  h = lrelu(zh)
    float* dZh = H + BP * LP;                   // [BP][LP]
    float* part = dZh + BP * LP;                // [NWARP][BP][LP] per-warp sums over the warp's ETF columns
    int* rows = reinterpret_cast<int*>(part + (size_t)NWARP * BP * LP);   // [BP]
    float* red = reinterpret_cast<float*>(rows + BP);                     // [32]
    // LP = 32: the 32 gradient accumulators of dWd and the 32 values of the dzh reduction do not fit the 128 registers of
    // a 512-thread CTA together, so dzy is parked in shared memory (a private column per thread) and dWd gets its own pass
    constexpr bool SPLIT = LP * EPT > 16;
    float* dzyS = red + 32;                                               // [BP][NT*EPT] (SPLIT only)

    const size_t EL = (size_t)E * L;
    const float* X = p.X + (size_t)(p.task_x ? p.task_x[task] : 0) * p.T * E;
    const float* w0 = p.W0 + (size_t)task * 2 * EL;
    const int e0 = tid * EPT;
    float2 We[EPT][LH], Wd[LH][EPT];            // We[k][j] = We[e0+k][2j, 2j+1];  Wd[j][k] = Wd[2j, 2j+1][e0+k]
    auto w_at = [&](size_t base, int idx, bool in) { return in ? w0[base + idx] : 0.f; };
#pragma unroll
    for (int k = 0; k < EPT; ++k)
#pragma unroll
        for (int j = 0; j < LH; ++j) {
            const bool ine = e0 + k < E;
            We[k][j] = make_float2(w_at(0, (e0 + k) * L + 2 * j, ine && 2 * j < L), w_at(0, (e0 + k) * L + 2 * j + 1, ine && 2 * j + 1 < L));
            Wd[j][k] = make_float2(w_at(EL, 2 * j * E + e0 + k, ine && 2 * j < L), w_at(EL, (2 * j + 1) * E + e0 + k, ine && 2 * j + 1 < L));
        }
    // Nadam state: slot s of thread t at mv[s*NT + t] (m) and mv[(NS + s)*NT + t] (v); slots = We then Wd
    constexpr int NS = 2 * EPT * LP;
    // (in shared memory when it fits beside the tiles — LP <= 16: 64 / 128 KB — else in the caller's workspace, L2-resident)
    float* mv = MVS ? dzyS + (SPLIT ? (size_t)BP * NT * EPT : 0) : p.ws + (size_t)task * (2 * NS) * NT;
    for (int s = 0; s < 2 * NS; ++s) mv[(size_t)s * NT + tid] = 0.f;

    const int* perm_t = p.perm ? p.perm + (size_t)task * p.perm_task_stride : nullptr;
    const float alpha = p.alpha, omb1 = 1.f - p.b1, omb2 = 1.f - p.b2;
    const int n_fit = p.n_fit, n_val = p.n_val;
    int step = 0, wait = 0, epochs_run = 0, st = 0;
    float best = INFINITY;

    auto load_x = [&](int r, int nb, float (&x)[EPT]) {          // the thread's columns of batch row r (zeros past the batch / E)
        const float* xr = X + (size_t)rows[r < nb ? r : 0] * E + e0;
#pragma unroll
        for (int k = 0; k < EPT; ++k) x[k] = (r < nb && e0 + k < E) ? __ldg(xr + k) : 0.f;
    };
    auto cross_warp = [&](int o) -> float {                      // sum of the warps' sums for output o, in warp order
        float sacc = 0.f;
#pragma unroll
        for (int w = 0; w < 16 / EPT; ++w)                        // (independent loads: unrolled over the most warps a CTA has)
            if (w < NWARP) sacc += part[(size_t)w * BP * LP + o];
        return sacc;
    };
    const float4* sched = p.sched;
    float mu1 = 0.f;
    float3 rc = make_float3(1.f, 1.f, 1.f);
    const NadamK nk{p.b1, omb1, p.b2, omb2, p.lr, p.eps};
    auto upd = [&](float& w, float g, int slot, const float4& sa) {
        float* mp = mv + (size_t)slot * NT + tid;
        float* vp = mv + (size_t)(NS + slot) * NT + tid;
        float mm = *mp, vv = *vp;
        w = nadam_step(w, g, mm, vv, sa, mu1, rc, nk);
        *mp = mm; *vp = vv;
    };

    auto run_batch = [&](int nb, bool train) -> float {
        // ---- pass 1: zh = x We (reduce over ETFs)
        for (int r0 = 0; r0 < nb; r0 += RG) {
            float v[32];
#pragma unroll
            for (int g = 0; g < RG; ++g) {
                float x[EPT];
                load_x(r0 + g, nb, x);
#pragma unroll
                for (int j = 0; j < LH; ++j) {
                    float2 a = make_float2(0.f, 0.f);
#pragma unroll
                    for (int k = 0; k < EPT; ++k) a = __ffma2_rn(make_float2(x[k], x[k]), We[k][j], a);
                    v[g * LP + 2 * j] = a.x; v[g * LP + 2 * j + 1] = a.y;
                }
            }
            int j0 = 0;
            const float sum = warp_scatter_sum<32, 16>(v, lane, j0);
            part[(size_t)warp * BP * LP + r0 * LP + j0] = sum;
        }
        __syncthreads();
        for (int o = tid; o < nb * LP; o += NT) H[o] = lrelu(cross_warp(o), alpha);
        __syncthreads();
        // ---- pass 2: zy = h Wd, error, dzy; dWd += h^T dzy; dzh partials = dzy Wd^T (reduce over ETFs)
        const float coef = 2.f / (float)(nb * E);
        float sse = 0.f;
        float2 gWd[LH][EPT];
#pragma unroll
        for (int j = 0; j < LH; ++j)
#pragma unroll
            for (int k = 0; k < EPT; ++k) gWd[j][k] = make_float2(0.f, 0.f);
        for (int r0 = 0; r0 < nb; r0 += RG) {
            float v[32];
#pragma unroll
            for (int g = 0; g < RG; ++g) {
                const int r = r0 + g;
                float x[EPT], dzy[EPT];
                float2 h[LH];
                load_x(r, nb, x);
#pragma unroll
                for (int l4 = 0; l4 < LP / 4; ++l4) {
                    const float4 hv = *reinterpret_cast<const float4*>(H + (r < nb ? r : 0) * LP + 4 * l4);
                    h[2 * l4] = make_float2(hv.x, hv.y); h[2 * l4 + 1] = make_float2(hv.z, hv.w);
                }
#pragma unroll
                for (int k = 0; k < EPT; ++k) {
                    float2 a = make_float2(0.f, 0.f);
#pragma unroll
                    for (int j = 0; j < LH; ++j) a = __ffma2_rn(h[j], Wd[j][k], a);
                    const float zy = a.x + a.y;
                    const float d = r < nb ? lrelu(zy, alpha) - x[k] : 0.f;       // columns past E: zy = x = 0
                    sse = fmaf(d, d, sse);
                    dzy[k] = lrelu_grad(zy, coef * d, alpha);
                }
                if (train) {
                    if constexpr (SPLIT) {
#pragma unroll
                        for (int k = 0; k < EPT; ++k) dzyS[(size_t)(r < nb ? r : BP - 1) * NT * EPT + tid * EPT + k] = dzy[k];
                    }
#pragma unroll
                    for (int j = 0; j < LH; ++j) {
                        float2 a = make_float2(0.f, 0.f);
#pragma unroll
                        for (int k = 0; k < EPT; ++k) {
                            const float2 dd = make_float2(dzy[k], dzy[k]);
                            if constexpr (!SPLIT) gWd[j][k] = __ffma2_rn(h[j], dd, gWd[j][k]);
                            a = __ffma2_rn(dd, Wd[j][k], a);
                        }
                        v[g * LP + 2 * j] = a.x; v[g * LP + 2 * j + 1] = a.y;
                    }
                }
            }
            if (train) {
                int j0 = 0;
                const float sum = warp_scatter_sum<32, 16>(v, lane, j0);
                part[(size_t)warp * BP * LP + r0 * LP + j0] = sum;
            }
        }
        if (!train) return sse;
        if constexpr (SPLIT) {                  // dWd += h^T dzy from the parked dzy column (the thread's own: no barrier)
            for (int r = 0; r < nb; ++r) {
                float2 dd[EPT];
#pragma unroll
                for (int k = 0; k < EPT; ++k) { const float d = dzyS[(size_t)r * NT * EPT + tid * EPT + k]; dd[k] = make_float2(d, d); }
#pragma unroll
                for (int l4 = 0; l4 < LP / 4; ++l4) {
                    const float4 hv = *reinterpret_cast<const float4*>(H + r * LP + 4 * l4);
#pragma unroll
                    for (int k = 0; k < EPT; ++k) {
                        gWd[2 * l4][k] = __ffma2_rn(make_float2(hv.x, hv.y), dd[k], gWd[2 * l4][k]);
                        gWd[2 * l4 + 1][k] = __ffma2_rn(make_float2(hv.z, hv.w), dd[k], gWd[2 * l4 + 1][k]);
                    }
                }
            }
        }
        // ---- Nadam on the thread's own weights (Wd now: its gradient is complete and pass 3 does not read Wd)
        const float4 sc = __ldg(sched + 2 * step);
        mu1 = __ldg(sched + 2 * step + 1).x;
        rc = make_float3(rcp_refined(sc.x), rcp_refined(sc.y), rcp_refined(sc.z));
#pragma unroll
        for (int j = 0; j < LH; ++j)
#pragma unroll
            for (int k = 0; k < EPT; ++k) {
                upd(Wd[j][k].x, gWd[j][k].x, EPT * LP + (2 * j) * EPT + k, sc);
                upd(Wd[j][k].y, gWd[j][k].y, EPT * LP + (2 * j + 1) * EPT + k, sc);
            }
        __syncthreads();
        for (int o = tid; o < nb * LP; o += NT) dZh[o] = lrelu_grad(H[o], cross_warp(o), alpha);   // sign(h) == sign(zh)
        __syncthreads();
        // ---- pass 3: dWe = x^T dzh
        float2 gWe[EPT][LH];
#pragma unroll
        for (int k = 0; k < EPT; ++k)
#pragma unroll
            for (int j = 0; j < LH; ++j) gWe[k][j] = make_float2(0.f, 0.f);
#pragma unroll 2
        for (int r = 0; r < nb; ++r) {
            float x[EPT];
            load_x(r, nb, x);
#pragma unroll
            for (int l4 = 0; l4 < LP / 4; ++l4) {
                const float4 gv = *reinterpret_cast<const float4*>(dZh + r * LP + 4 * l4);
#pragma unroll
                for (int k = 0; k < EPT; ++k) {
                    const float2 xx = make_float2(x[k], x[k]);
                    gWe[k][2 * l4] = __ffma2_rn(xx, make_float2(gv.x, gv.y), gWe[k][2 * l4]);
                    gWe[k][2 * l4 + 1] = __ffma2_rn(xx, make_float2(gv.z, gv.w), gWe[k][2 * l4 + 1]);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < EPT; ++k)
#pragma unroll
            for (int j = 0; j < LH; ++j) {
                upd(We[k][j].x, gWe[k][j].x, k * LP + 2 * j, sc);
                upd(We[k][j].y, gWe[k][j].y, k * LP + 2 * j + 1, sc);
            }
        ++step;
        return sse;
    };

    for (int epoch = 0; epoch < p.max_epochs; ++epoch) {
        float lsum = 0.f, vsum = 0.f;
        for (int lo = 0; lo < n_fit; lo += B) {
            const int nb = min(B, n_fit - lo);
            __syncthreads();
            for (int q = tid; q < nb; q += NT) rows[q] = perm_t ? __ldg(perm_t + (size_t)epoch * n_fit + lo + q) : lo + q;
            __syncthreads();
            lsum += run_batch(nb, true);
        }
        for (int lo = 0; lo < n_val; lo += B) {
            const int nb = min(B, n_val - lo);
            __syncthreads();
            for (int q = tid; q < nb; q += NT) rows[q] = n_fit + lo + q;
            __syncthreads();
            vsum += run_batch(nb, false);
        }
        // epoch losses: fixed-order block sums
        const float l1 = warp_sum(lsum), v1 = warp_sum(vsum);
        __syncthreads();
        if (lane == 0) { red[warp] = l1; red[16 + warp] = v1; }
        __syncthreads();
        float loss = 0.f, val = 0.f;
        for (int w = 0; w < NWARP; ++w) { loss += red[w]; val += red[16 + w]; }
        loss = loss / ((float)E * n_fit);
        val = n_val > 0 ? val / ((float)E * n_val) : NAN;
        if (tid == 0) {
            p.loss_hist[(size_t)task * p.max_epochs + epoch] = loss;
            p.val_hist[(size_t)task * p.max_epochs + epoch] = val;
        }
        epochs_run = epoch + 1;
        if (!isfinite(loss)) st = 1;
        if (p.early_stop && n_val > 0) {         // no validation rows: no val_loss, Keras never stops (see fit_kernel.cu)
            ++wait;
            if (val < best) { best = val; wait = 0; }
            if (wait >= p.patience && epoch > 0) break;
        }
    }
    float* wo = p.W_out + (size_t)task * 2 * EL;
#pragma unroll
    for (int k = 0; k < EPT; ++k)
#pragma unroll
        for (int j = 0; j < LH; ++j)
            if (e0 + k < E) {
                if (2 * j < L) { wo[(size_t)(e0 + k) * L + 2 * j] = We[k][j].x; wo[EL + (size_t)(2 * j) * E + e0 + k] = Wd[j][k].x; }
                if (2 * j + 1 < L) { wo[(size_t)(e0 + k) * L + 2 * j + 1] = We[k][j].y; wo[EL + (size_t)(2 * j + 1) * E + e0 + k] = Wd[j][k].y; }
            }
    if (tid == 0) { p.epochs_run[task] = epochs_run; p.status[task] = st; }
}

// ------------------------------------------------------------------------------------------------ wide fit, v3
// v2 spends two fifths of its instructions on the shuffle scatter-reductions of the two contractions that sum over the
// ETFs (zh = x We, dzh = dzy Wd^T).  v3 (LP = 8: configs[3]'s width) keeps v2's ownership for everything thread-local
// — thread t owns ETF columns 2t, 2t+1: its weights in registers, zy / dWd / dWe / Nadam as before — and runs those two
// contractions as ROW-mapped register tiles instead: the mini-batch sits in a shared tile [48][516] (cp.async; 516 = 4
// mod 32 words, so eight lanes reading eight different rows with LDS.128 touch eight different bank groups), shared copies
// of We [e][8] and Wd^T [e][8] are rewritten by their owners after every update, thread (row triple, 32-ETF chunk)
// accumulates 3 x 8 partial sums over its chunk (per 4 ETFs: 3 LDS.128 of x, 8 broadcast LDS.128 of weights, 48 FFMA2)
// and the 16 chunks' partial sums meet in shared memory in chunk order.  Pass 2 overwrites the tile's x with dzy in place
// (each thread the elements it just read), so dzh's pass reads it back row-mapped; dWe re-reads x from global memory
// (L1/L2, coalesced) as in v2.  Nadam's m, v sit in shared memory.  220 KB of shared memory, one CTA of 256 threads per SM.
constexpr int W3_TS = 516;          // tile row stride (floats)
constexpr int W3_EP = 512;          // ETF capacity
constexpr int W3_BP = 48;           // batch rows (16 row triples)
constexpr int W3_NCH = 16;          // 32-ETF chunks whose partial sums meet in shared memory

__device__ __forceinline__ void cp_async16(float* dst_smem, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// EPT = ETF columns per thread: 2 (256 threads) or 1 (512 threads: twice the warps to hide the dependent-issue latency of
// the short chains, for about a sixth more instructions — every thread loads the h / dzh rows it multiplies with)
template <int EPT>
__global__ void __launch_bounds__(512 / EPT) fit_wide3_kernel(const WideFitParams p) {
    extern __shared__ __align__(16) float sm[];
    constexpr int LP = 8, LH = 4, NT = 512 / EPT, BP = W3_BP, TS = W3_TS;
    constexpr int CW = 16 * EPT;                    // ETFs of a thread's chunk in the row-mapped contractions
    constexpr int NS = 2 * EPT * LP;                                   // weight slots per thread
    const int task = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int E = p.E, L = p.L, B = p.batch;
    float* H = sm;                                  // [BP][LP]
    float* dZh = H + BP * LP;                       // [BP][LP]
    float* part = dZh + BP * LP;                    // [NCH][BP][LP]
    int* rows2 = reinterpret_cast<int*>(part + W3_NCH * BP * LP);      // [2][BP] rows of this batch and of the next
    float* red = reinterpret_cast<float*>(rows2 + 2 * BP);             // [32]
    float* tile = red + 32;                         // [BP][TS]   x of the mini-batch, then dzy
    float* WeS = tile + BP * TS;                    // [EP][LP]   We rows
    float* WdT = WeS + W3_EP * LP;                  // [EP][LP]   Wd columns
    float* mv = WdT + W3_EP * LP;                   // [2*NS][NT] Nadam m then v, slot-major

    const size_t EL = (size_t)E * L;
    const float* X = p.X + (size_t)(p.task_x ? p.task_x[task] : 0) * p.T * E;
    const float* w0 = p.W0 + (size_t)task * 2 * EL;
    const int e0 = tid * EPT;
    float2 We[EPT][LH], Wd[LH][EPT];
    auto w_at = [&](size_t base, int idx, bool in) { return in ? w0[base + idx] : 0.f; };
#pragma unroll
    for (int k = 0; k < EPT; ++k)
#pragma unroll
        for (int j = 0; j < LH; ++j) {
            const bool ine = e0 + k < E;
            We[k][j] = make_float2(w_at(0, (e0 + k) * L + 2 * j, ine && 2 * j < L), w_at(0, (e0 + k) * L + 2 * j + 1, ine && 2 * j + 1 < L));
            Wd[j][k] = make_float2(w_at(EL, 2 * j * E + e0 + k, ine && 2 * j < L), w_at(EL, (2 * j + 1) * E + e0 + k, ine && 2 * j + 1 < L));
        }
    auto publish_We = [&]() {
#pragma unroll
        for (int k = 0; k < EPT; ++k) {
            float4* d = reinterpret_cast<float4*>(WeS + (e0 + k) * LP);
            d[0] = make_float4(We[k][0].x, We[k][0].y, We[k][1].x, We[k][1].y);
            d[1] = make_float4(We[k][2].x, We[k][2].y, We[k][3].x, We[k][3].y);
        }
    };
    auto publish_Wd = [&]() {
#pragma unroll
        for (int k = 0; k < EPT; ++k) {
            float4* d = reinterpret_cast<float4*>(WdT + (e0 + k) * LP);
            d[0] = make_float4(Wd[0][k].x, Wd[0][k].y, Wd[1][k].x, Wd[1][k].y);
            d[1] = make_float4(Wd[2][k].x, Wd[2][k].y, Wd[3][k].x, Wd[3][k].y);
        }
    };
    publish_We();
    publish_Wd();
    for (int s = 0; s < 2 * NS; ++s) mv[s * NT + tid] = 0.f;
    for (int i = tid; i < BP * TS; i += NT) tile[i] = 0.f;              // columns past E and rows past the batch stay zero

    const int* perm_t = p.perm ? p.perm + (size_t)task * p.perm_task_stride : nullptr;
    const float alpha = p.alpha, omb1 = 1.f - p.b1, omb2 = 1.f - p.b2;
    const int n_fit = p.n_fit, n_val = p.n_val;
    int step = 0, wait = 0, epochs_run = 0, st = 0;
    float best = INFINITY;
    const float4* sched = p.sched;
    float mu1 = 0.f;
    float3 rc = make_float3(1.f, 1.f, 1.f);
    const NadamK nk{p.b1, omb1, p.b2, omb2, p.lr, p.eps};
    auto upd = [&](float& w, float g, int slot, const float4& sa) {
        float* mp = mv + slot * NT + tid;
        float* vp = mv + (NS + slot) * NT + tid;
        float mm = *mp, vv = *vp;
        w = nadam_step(w, g, mm, vv, sa, mu1, rc, nk);
        *mp = mm; *vp = vv;
    };
    const bool vec_rows = (E & 3) == 0;             // 16-byte row pieces (rows of X start 16-byte aligned)
    const int E4 = (E + 3) >> 2;

    // a mini-batch's rows into the tile (rows past nb: zeros); asynchronous: cp_async_wait_all + barrier before use
    auto fill_tile = [&](const int* rows, int nb) {
        for (int r = warp; r < BP; r += NT / 32) {                     // a warp per row, lanes over its 16-byte pieces
            float* d0 = tile + r * TS;
            if (r < nb) {
                const float* src = X + (size_t)rows[r] * E;
                for (int c4 = lane; c4 < E4; c4 += 32) {
                    if (vec_rows) cp_async16(d0 + 4 * c4, src + 4 * c4);
                    else {
#pragma unroll
                        for (int q = 0; q < 4; ++q) d0[4 * c4 + q] = 4 * c4 + q < E ? __ldg(src + 4 * c4 + q) : 0.f;
                    }
                }
            } else {
                for (int c4 = lane; c4 < E4; c4 += 32) *reinterpret_cast<float4*>(d0 + 4 * c4) = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    };
    // part[ch][r][l] = sum over chunk ch's ETFs of tile[r][e] * Ws[e][l]  for the thread's row triple
    const int rt = tid & 15, ch = tid >> 4;          // row triple, chunk of CW ETFs (EPT = 1: 32 chunks, lane pairs l, l^16 add up)
    auto row_contract = [&](const float* Ws) {
        float2 acc[3][LH];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < LH; ++j) acc[i][j] = make_float2(0.f, 0.f);
        const float* t0 = tile + (3 * rt) * TS + CW * ch;
        const float* wb = Ws + (CW * ch) * LP;
#pragma unroll 2
        for (int e4 = 0; e4 < CW / 4; ++e4) {
            float4 xv[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) xv[i] = *reinterpret_cast<const float4*>(t0 + i * TS + 4 * e4);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 wa = *reinterpret_cast<const float4*>(wb + (4 * e4 + q) * LP);
                const float4 wc = *reinterpret_cast<const float4*>(wb + (4 * e4 + q) * LP + 4);
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    const float xe = q == 0 ? xv[i].x : (q == 1 ? xv[i].y : (q == 2 ? xv[i].z : xv[i].w));
                    const float2 xx = make_float2(xe, xe);
                    acc[i][0] = __ffma2_rn(xx, make_float2(wa.x, wa.y), acc[i][0]);
                    acc[i][1] = __ffma2_rn(xx, make_float2(wa.z, wa.w), acc[i][1]);
                    acc[i][2] = __ffma2_rn(xx, make_float2(wc.x, wc.y), acc[i][2]);
                    acc[i][3] = __ffma2_rn(xx, make_float2(wc.z, wc.w), acc[i][3]);
                }
            }
        }
        if constexpr (EPT == 1) {                    // the two 16-ETF chunks of a warp: lower chunk + upper chunk
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < LH; ++j) {
                    const float ox = __shfl_xor_sync(0xffffffffu, acc[i][j].x, 16), oy = __shfl_xor_sync(0xffffffffu, acc[i][j].y, 16);
                    acc[i][j] = (lane & 16) ? make_float2(ox + acc[i][j].x, oy + acc[i][j].y) : make_float2(acc[i][j].x + ox, acc[i][j].y + oy);
                }
        }
        if (EPT == 2 || lane < 16) {
            const int pc = EPT == 2 ? ch : (ch >> 1);
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                float4* d = reinterpret_cast<float4*>(part + (pc * BP + 3 * rt + i) * LP);
                d[0] = make_float4(acc[i][0].x, acc[i][0].y, acc[i][1].x, acc[i][1].y);
                d[1] = make_float4(acc[i][2].x, acc[i][2].y, acc[i][3].x, acc[i][3].y);
            }
        }
    };
    auto chunk_sum = [&](int o) -> float {           // the 16 chunks' sums for output o, in chunk order
        float sacc = 0.f;
#pragma unroll
        for (int c = 0; c < W3_NCH; ++c) sacc += part[c * BP * LP + o];
        return sacc;
    };

    // Batches of an epoch in order: the training mini-batches, then the validation chunks.  The tile of batch b + 1 is
    // requested as soon as batch b has no more use for the tile (after dzh's pass; after pass 2 of a validation chunk),
    // so the copies run under the Nadam updates and dWe's pass.
    const int ntb = (n_fit + B - 1) / B, nvb = (n_val + B - 1) / B, nbt = ntb + nvb;
    auto batch_of = [&](int b, int& lo, int& nb, bool& train) {       // b in [0, nbt)
        train = b < ntb;
        lo = train ? b * B : n_fit + (b - ntb) * B;
        nb = train ? min(B, n_fit - b * B) : min(B, n_val - (b - ntb) * B);
    };
    auto write_rows = [&](int* rows, int epoch, int b) {              // (made visible by the next barrier)
        int lo, nb; bool train;
        batch_of(b, lo, nb, train);
        for (int q = tid; q < nb; q += NT)
            rows[q] = (train && perm_t) ? __ldg(perm_t + (size_t)epoch * n_fit + lo + q) : lo + q;
    };
    auto prefetch_next = [&](int epoch, int b) {                      // the tile is free: request batch b + 1 (or the next epoch's first)
        int e2 = epoch, b2 = b + 1;
        if (b2 == nbt) { b2 = 0; ++e2; }
        if (e2 >= p.max_epochs) return;
        int lo, nb; bool train;
        batch_of(b2, lo, nb, train);
        fill_tile(rows2 + ((e2 * nbt + b2) & 1) * BP, nb);
    };

    auto run_batch = [&](int epoch, int b) -> float {
        int lo_, nb; bool train;
        batch_of(b, lo_, nb, train);
        const int* rows = rows2 + ((epoch * nbt + b) & 1) * BP;
        {   // the next batch's rows (read by prefetch_next after at least one barrier)
            int e2 = epoch, b2 = b + 1;
            if (b2 == nbt) { b2 = 0; ++e2; }
            if (e2 < p.max_epochs) write_rows(rows2 + ((e2 * nbt + b2) & 1) * BP, e2, b2);
        }
        cp_async_wait_all();
        __syncthreads();
        // ---- pass 1: zh = x We (row-mapped), h
        row_contract(WeS);
        __syncthreads();
        for (int o = tid; o < BP * LP; o += NT) H[o] = lrelu(chunk_sum(o), alpha);
        __syncthreads();
        // ---- pass 2 (owner-mapped): zy = h Wd, error, dzy -> tile in place; dWd += h^T dzy
        const float coef = 2.f / (float)(nb * E);
        float sse = 0.f;
        float2 gWd[LH][EPT];
#pragma unroll
        for (int j = 0; j < LH; ++j)
#pragma unroll
            for (int k = 0; k < EPT; ++k) gWd[j][k] = make_float2(0.f, 0.f);
#pragma unroll 2
        for (int r = 0; r < nb; ++r) {
            const float4 h0 = *reinterpret_cast<const float4*>(H + r * LP), h1 = *reinterpret_cast<const float4*>(H + r * LP + 4);
            const float2 h[LH] = {make_float2(h0.x, h0.y), make_float2(h0.z, h0.w), make_float2(h1.x, h1.y), make_float2(h1.z, h1.w)};
            float* tp = tile + r * TS + e0;
            float x[EPT], dzy[EPT];
            if constexpr (EPT == 2) { const float2 xv = *reinterpret_cast<const float2*>(tp); x[0] = xv.x; x[1] = xv.y; }
            else x[0] = *tp;
#pragma unroll
            for (int k = 0; k < EPT; ++k) {
                float2 a = make_float2(0.f, 0.f);
#pragma unroll
                for (int j = 0; j < LH; ++j) a = __ffma2_rn(h[j], Wd[j][k], a);
                const float zy = a.x + a.y;
                const float d = lrelu(zy, alpha) - x[k];                 // columns past E: zy = x = 0
                sse = fmaf(d, d, sse);
                dzy[k] = lrelu_grad(zy, coef * d, alpha);
            }
            if (train) {
                if constexpr (EPT == 2) *reinterpret_cast<float2*>(tp) = make_float2(dzy[0], dzy[1]);
                else *tp = dzy[0];
#pragma unroll
                for (int j = 0; j < LH; ++j)
#pragma unroll
                    for (int k = 0; k < EPT; ++k) gWd[j][k] = __ffma2_rn(h[j], make_float2(dzy[k], dzy[k]), gWd[j][k]);
            }
        }
        if (!train) { __syncthreads(); prefetch_next(epoch, b); return sse; }
        __syncthreads();
        // ---- pass 2b: dzh = (dzy Wd^T) o lrelu'(zh), row-mapped, with the forward pass's Wd
        row_contract(WdT);
        __syncthreads();
        prefetch_next(epoch, b);
        for (int o = tid; o < BP * LP; o += NT) dZh[o] = lrelu_grad(H[o], chunk_sum(o), alpha);   // sign(h) == sign(zh)
        // ---- Nadam on Wd (its gradient is complete; nothing reads WdT before the next step's pass 2b)
        const float4 sc = __ldg(sched + 2 * step);
        mu1 = __ldg(sched + 2 * step + 1).x;
        rc = make_float3(rcp_refined(sc.x), rcp_refined(sc.y), rcp_refined(sc.z));
#pragma unroll
        for (int j = 0; j < LH; ++j)
#pragma unroll
            for (int k = 0; k < EPT; ++k) {
                upd(Wd[j][k].x, gWd[j][k].x, EPT * LP + (2 * j) * EPT + k, sc);
                upd(Wd[j][k].y, gWd[j][k].y, EPT * LP + (2 * j + 1) * EPT + k, sc);
            }
        publish_Wd();
        __syncthreads();
        // ---- pass 3 (owner-mapped): dWe = x^T dzh, x from global memory (the tile holds dzy now)
        float2 gWe[EPT][LH];
#pragma unroll
        for (int k = 0; k < EPT; ++k)
#pragma unroll
            for (int j = 0; j < LH; ++j) gWe[k][j] = make_float2(0.f, 0.f);
#pragma unroll 8
        for (int r = 0; r < nb; ++r) {
            const float* xr = X + (size_t)rows[r] * E + e0;
            float x[EPT];
#pragma unroll
            for (int k = 0; k < EPT; ++k) x[k] = e0 + k < E ? __ldg(xr + k) : 0.f;
            const float4 g0 = *reinterpret_cast<const float4*>(dZh + r * LP), g1 = *reinterpret_cast<const float4*>(dZh + r * LP + 4);
#pragma unroll
            for (int k = 0; k < EPT; ++k) {
                const float2 xx = make_float2(x[k], x[k]);
                gWe[k][0] = __ffma2_rn(xx, make_float2(g0.x, g0.y), gWe[k][0]);
                gWe[k][1] = __ffma2_rn(xx, make_float2(g0.z, g0.w), gWe[k][1]);
                gWe[k][2] = __ffma2_rn(xx, make_float2(g1.x, g1.y), gWe[k][2]);
                gWe[k][3] = __ffma2_rn(xx, make_float2(g1.z, g1.w), gWe[k][3]);
            }
        }
#pragma unroll
        for (int k = 0; k < EPT; ++k)
#pragma unroll
            for (int j = 0; j < LH; ++j) {
                upd(We[k][j].x, gWe[k][j].x, k * LP + 2 * j, sc);
                upd(We[k][j].y, gWe[k][j].y, k * LP + 2 * j + 1, sc);
            }
        publish_We();
        ++step;
        __syncthreads();
        return sse;
    };

    write_rows(rows2, 0, 0);
    __syncthreads();
    {
        int lo, nb; bool train;
        batch_of(0, lo, nb, train);
        fill_tile(rows2, nb);
    }
    for (int epoch = 0; epoch < p.max_epochs; ++epoch) {
        float lsum = 0.f, vsum = 0.f;
        for (int b = 0; b < ntb; ++b) lsum += run_batch(epoch, b);
        for (int b = ntb; b < nbt; ++b) vsum += run_batch(epoch, b);
        const float l1 = warp_sum(lsum), v1 = warp_sum(vsum);
        if (lane == 0) { red[warp] = l1; red[16 + warp] = v1; }
        __syncthreads();
        float loss = 0.f, val = 0.f;
        for (int w = 0; w < NT / 32; ++w) { loss += red[w]; val += red[16 + w]; }
        __syncthreads();
        loss = loss / ((float)E * n_fit);
        val = n_val > 0 ? val / ((float)E * n_val) : NAN;
        if (tid == 0) {
            p.loss_hist[(size_t)task * p.max_epochs + epoch] = loss;
            p.val_hist[(size_t)task * p.max_epochs + epoch] = val;
        }
        epochs_run = epoch + 1;
        if (!isfinite(loss)) st = 1;
        if (p.early_stop && n_val > 0) {         // no validation rows: no val_loss, Keras never stops (see fit_kernel.cu)
            ++wait;
            if (val < best) { best = val; wait = 0; }
            if (wait >= p.patience && epoch > 0) break;
        }
    }
    cp_async_wait_all();                             // (a requested tile may still be in flight after an early stop)
    float* wo = p.W_out + (size_t)task * 2 * EL;
#pragma unroll
    for (int k = 0; k < EPT; ++k)
#pragma unroll
        for (int j = 0; j < LH; ++j)
            if (e0 + k < E) {
                if (2 * j < L) { wo[(size_t)(e0 + k) * L + 2 * j] = We[k][j].x; wo[EL + (size_t)(2 * j) * E + e0 + k] = Wd[j][k].x; }
                if (2 * j + 1 < L) { wo[(size_t)(e0 + k) * L + 2 * j + 1] = We[k][j].y; wo[EL + (size_t)(2 * j + 1) * E + e0 + k] = Wd[j][k].y; }
            }
    if (tid == 0) { p.epochs_run[task] = epochs_run; p.status[task] = st; }
}

// ------------------------------------------------------------------------------------------------ wide evaluation
struct WideEvalParams {
    const double* etf; const double* hfd; const double* rf;   // full histories [T][E], [T][F], [T]
    const int* task_row0;      // first row of each task's (window+2)-row test block
    const float* W;            // [n_tasks][2*E*L] trained kernels
    double* replica;           // [n_tasks][F]   replica return of row row0+window+1
    int* status;
    int n_tasks, E, F, L, window;
    float alpha;
};

// One CTA per task.  Latent rows, mask, c_l = sum_e etf[t+1][e] mask_e Wd[l][e], s_l = sum_e mask_e Wd[l][e] once;
// then per fund (threads over funds): Householder-QR least squares (the pinv-OLS of :151 for full-rank windows),
// helper.normalization, and replica = (1 - norm*beta.s) rf + norm*beta.c  (:167-168, :188-190 with the sums over
// ETFs regrouped through the latent factors, so a fund costs O(L*window) instead of O(E)).
__global__ void __launch_bounds__(WIDE_THREADS) eval_wide_kernel(const WideEvalParams p) {
    extern __shared__ __align__(16) unsigned char smraw[];
    const int task = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int E = p.E, F = p.F, L = p.L, window = p.window, R = window + 2;
    double* Q = reinterpret_cast<double*>(smraw);      // [window][L] : R factor in the upper triangle after QR
    double* V = Q + window * L;                         // [L][window] Householder vectors
    double* bh = V + L * window;                        // [L] 2/|v|^2
    double* cl = bh + L;                                // [L]
    double* sl = cl + L;                                // [L]
    float* M = reinterpret_cast<float*>(sl + L);        // [R][L] latent factors of the test block
    unsigned char* mask = reinterpret_cast<unsigned char*>(M + R * L);   // [E]
    __shared__ int s_status;
    const size_t EL = (size_t)E * L;
    const float* We = p.W + (size_t)task * 2 * EL;
    const float* Wd = We + EL;
    const int row0 = p.task_row0[task];
    const double* X = p.etf + (size_t)row0 * E;
    const double* Y = p.hfd + (size_t)row0 * F;
    if (tid == 0) s_status = 0;

    // latent rows, float32 sequential (encoder.predict on raw rows, :140)
    for (int o = tid; o < R * L; o += WIDE_THREADS) {
        const int r = o / L, l = o - r * L;
        float acc = 0.f;
        for (int e = 0; e < E; ++e) acc = fmaf((float)X[(size_t)r * E + e], We[(size_t)e * L + l], acc);
        M[o] = lrelu(acc, p.alpha);
    }
    __syncthreads();
    // mask from the row after the OLS window (:163-166), float32
    for (int e = tid; e < E; e += WIDE_THREADS) {
        float acc = 0.f;
        for (int l = 0; l < L; ++l) acc = fmaf(M[window * L + l], Wd[(size_t)l * E + e], acc);
        mask[e] = acc < 0.f ? 1 : 0;
    }
    // least squares on the latent block: QR of M (window x L) when L <= window; when L > window the system is
    // underdetermined and pinv (:151) returns the minimum-norm solution, obtained from the QR of M^T (L x window)
    const bool minnorm = L > window;
    const int nr = minnorm ? L : window, nc = minnorm ? window : L;
    for (int o = tid; o < window * L; o += WIDE_THREADS) {
        const int r = o / L, l = o - r * L;               // M[r][l]
        if (minnorm) Q[l * nc + r] = (double)M[o]; else Q[o] = (double)M[o];
    }
    __syncthreads();
    // c_l, s_l: a warp per latent factor
    for (int l = warp; l < L; l += WIDE_THREADS / 32) {
        double c = 0., s = 0.;
        for (int e = lane; e < E; e += 32) {
            const double w = (double)Wd[(size_t)l * E + e] * (mask[e] ? .2 : 1.);
            c += X[(size_t)(window + 1) * E + e] * w;
            s += w;
        }
        c = warp_sum(c); s = warp_sum(s);
        if (lane == 0) { cl[l] = c; sl[l] = s; }
    }
    // Householder QR (nr x nc), warp 0, lane = column
    if (warp == 0) {
        double dmax = 0., dmin = INFINITY;
        for (int k = 0; k < nc; ++k) {
            double sq = 0.;
            for (int r = k; r < nr; ++r) sq += Q[r * nc + k] * Q[r * nc + k];
            const double akk = Q[k * nc + k];
            const double nrm = sqrt(sq);
            const double alph = akk > 0. ? -nrm : nrm;
            const double vtv = sq - 2. * akk * alph + alph * alph;
            dmax = fmax(dmax, fabs(alph)); dmin = fmin(dmin, fabs(alph));
            __syncwarp();
            for (int r = lane; r < nr; r += 32) V[k * nr + r] = r < k ? 0. : Q[r * nc + k] - (r == k ? alph : 0.);
            if (lane == 0) bh[k] = vtv > 0. ? 2. / vtv : 0.;
            __syncwarp();
            for (int c = k + 1 + lane; c < nc; c += 32) {
                double wj = 0.;
                for (int r = k; r < nr; ++r) wj += V[k * nr + r] * Q[r * nc + c];
                wj *= bh[k];
                for (int r = k; r < nr; ++r) Q[r * nc + c] -= V[k * nr + r] * wj;
            }
            __syncwarp();
            if (lane == 0) Q[k * nc + k] = alph;
            __syncwarp();
        }
        if (!(dmin > 1e-13 * dmax) && lane == 0) s_status = 2;
    }
    __syncthreads();
    // funds
    constexpr int MAXW = 64;
    const double rf_next = p.rf[row0 + window + 1];
    for (int f = tid; f < F; f += WIDE_THREADS) {
        double y[MAXW], beta[MAXW];
        double my = 0.;
        for (int r = 0; r < window; ++r) { y[r] = Y[(size_t)r * F + f]; my += y[r]; }
        my /= window;
        double vy = 0.;
        for (int r = 0; r < window; ++r) vy += (y[r] - my) * (y[r] - my) / (window - 1);
        if (!minnorm) {
            for (int k = 0; k < L; ++k) {                   // Q^T y
                double wj = 0.;
                for (int r = k; r < window; ++r) wj += V[k * nr + r] * y[r];
                wj *= bh[k];
                for (int r = k; r < window; ++r) y[r] -= V[k * nr + r] * wj;
            }
            for (int l = L - 1; l >= 0; --l) {              // back substitution
                double acc = y[l];
                for (int j = l + 1; j < L; ++j) acc -= Q[l * nc + j] * beta[j];
                beta[l] = acc / Q[l * nc + l];
            }
        } else {
            // M^T = H_0..H_{w-1} [R; 0]  =>  beta = H_0..H_{w-1} [R^-T y; 0]
            for (int i = 0; i < window; ++i) {              // forward substitution with R^T
                double acc = y[i];
                for (int j = 0; j < i; ++j) acc -= Q[j * nc + i] * beta[j];
                beta[i] = acc / Q[i * nc + i];
            }
            for (int l = window; l < L; ++l) beta[l] = 0.;
            for (int k = window - 1; k >= 0; --k) {
                double wj = 0.;
                for (int r = k; r < L; ++r) wj += V[k * nr + r] * beta[r];
                wj *= bh[k];
                for (int r = k; r < L; ++r) beta[r] -= V[k * nr + r] * wj;
            }
        }
        double mr = 0.;                                     // helper.normalization (helper.py:10-17)
        for (int r = 0; r < window; ++r) {
            double rh = 0.;
            for (int l = 0; l < L; ++l) rh += (double)M[r * L + l] * beta[l];
            y[r] = rh; mr += rh;
        }
        mr /= window;
        double vr = 0.;
        for (int r = 0; r < window; ++r) vr += (y[r] - mr) * (y[r] - mr) / (window - 1);
        const double norm = sqrt(vy) / sqrt(vr);
        double bs = 0., bc = 0.;
        for (int l = 0; l < L; ++l) { bs += beta[l] * sl[l]; bc += beta[l] * cl[l]; }
        p.replica[(size_t)task * F + f] = (1. - norm * bs) * rf_next + norm * bc;
    }
    if (tid == 0 && p.status) p.status[task] = s_status;
}

}  // namespace hfrep

using namespace hfrep;

// v2 geometry: padded latent width, ETF columns per thread, threads per CTA (0: shape outside v2, use the v1 kernel)
struct Wide2Geom { int LP, EPT, NT; };
static Wide2Geom wide2_geom(int E, int L, int n_tasks) {
    Wide2Geom g{0, 0, 0};
    if (E > 512 || L > 32) return g;
    g.LP = L <= 8 ? 8 : (L <= 16 ? 16 : 32);
    // registers hold 2*EPT*LP weights per thread: 64 at most; few tasks -> more, narrower threads
    g.EPT = g.LP == 8 ? 2 : 1;                   // 16 weights of We and of Wd per thread (32 at L > 16): 256 or 512 threads
                                                 // (measured: 32 per thread halves the reduction work per FMA but leaves too
                                                 // few warps to hide the reduction's shuffle chain: 45.7 vs 29.4 ms at L = 8)
    g.NT = 32 * (((E + g.EPT - 1) / g.EPT + 31) / 32);
    return g;
}

extern "C" size_t hfrep_fit_wide_workspace(int n_tasks, int E, int L) {
    if (n_tasks <= 0 || E <= 0 || L <= 0) return 0;
    const Wide2Geom g = wide2_geom(E, L, n_tasks);
    const size_t v1 = 6 * (size_t)E * L, v2 = g.NT ? 4 * (size_t)g.EPT * g.LP * g.NT : 0;
    return sizeof(float) * (size_t)n_tasks * (v1 > v2 ? v1 : v2);
}

template <int LP, int EPT>
static int launch_fit_wide2(const WideFitParams& p, int NT, cudaStream_t stream) {
    const int BP = (p.batch + 3) & ~3, NWARP = NT / 32;
    const size_t smem = sizeof(float) * ((size_t)(2 + NWARP) * BP * LP + BP + 32 + (LP * EPT > 16 ? (size_t)BP * NT * EPT : 0));
    HFREP_REQUIRE(smem <= 227 * 1024, "fit_wide: %zu bytes of shared memory needed (batch too large)", smem);
    const size_t mv_bytes = sizeof(float) * 4 * EPT * LP * (size_t)NT;       // Nadam's m and v of the thread's 2*EPT*LP weights
    if (smem + mv_bytes <= 227 * 1024 && !getenv("HFREP_WIDE_MV_GLOBAL")) {
        auto kern = fit_wide2_kernel<LP, EPT, true>;
        HFREP_CUDA_OK(allow_max_dynamic_smem(kern));
        kern<<<p.n_tasks, NT, smem + mv_bytes, stream>>>(p);
    } else {
        auto kern = fit_wide2_kernel<LP, EPT, false>;
        HFREP_CUDA_OK(allow_max_dynamic_smem(kern));
        kern<<<p.n_tasks, NT, smem, stream>>>(p);
    }
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_fit_wide(const hfrep_fit_cfg* c, const float* X, int n_tasks, int L, const int32_t* task_x,
                              const float* W0, const int32_t* perm, int64_t perm_task_stride, const float* sched,
                              void* ws, size_t ws_bytes, float* W_out, float* loss_hist, float* val_hist,
                              int32_t* epochs_run, int32_t* status, void* stream) {
    HFREP_REQUIRE(c && X && W0 && sched && ws && W_out && loss_hist && val_hist && epochs_run && status,
                  "fit_wide: null argument");
    HFREP_REQUIRE(n_tasks >= 0, "fit_wide: n_tasks=%d", n_tasks);
    if (n_tasks == 0) return HFREP_OK;
    HFREP_REQUIRE(c->E >= 1 && c->E <= 4096, "fit_wide: E=%d outside 1..4096", c->E);
    HFREP_REQUIRE(L >= 1 && L <= WIDE_MAX_L, "fit_wide: L=%d outside 1..%d", L, WIDE_MAX_L);
    HFREP_REQUIRE(c->T >= 2 && c->n_fit >= 1 && c->n_fit <= c->T && c->batch >= 1, "fit_wide: T=%d n_fit=%d batch=%d", c->T,
                  c->n_fit, c->batch);
    HFREP_REQUIRE(ws_bytes >= hfrep_fit_wide_workspace(n_tasks, c->E, L), "fit_wide: workspace of %zu bytes needed, %zu given",
                  hfrep_fit_wide_workspace(n_tasks, c->E, L), ws_bytes);
    HFREP_REQUIRE(c->n_x == 1 || task_x, "fit_wide: n_x=%d needs task_x", c->n_x);
    HFREP_REQUIRE(((uintptr_t)sched & 15) == 0, "fit_wide: sched must be 16-byte aligned");
    WideFitParams p;
    p.X = X; p.task_x = task_x; p.W0 = W0; p.perm = perm; p.perm_task_stride = perm_task_stride;
    p.sched = reinterpret_cast<const float4*>(sched); p.ws = static_cast<float*>(ws); p.W_out = W_out;
    p.loss_hist = loss_hist; p.val_hist = val_hist; p.epochs_run = epochs_run; p.status = status;
    p.n_tasks = n_tasks; p.E = c->E; p.L = L; p.T = c->T; p.n_fit = c->n_fit; p.n_val = c->T - c->n_fit; p.batch = c->batch;
    p.max_epochs = c->max_epochs; p.patience = c->patience; p.early_stop = c->early_stop;
    p.alpha = c->alpha; p.lr = c->lr; p.b1 = c->beta_1; p.b2 = c->beta_2; p.eps = c->epsilon;
    const Wide2Geom g = getenv("HFREP_WIDE_V1") ? Wide2Geom{0, 0, 0} : wide2_geom(c->E, L, n_tasks);
    if (g.NT) {
        cudaStream_t st = (cudaStream_t)stream;
        if (g.LP == 8 && c->batch <= W3_BP && !getenv("HFREP_WIDE_V2")) {      // row-mapped reductions (v3)
            const int ept = getenv("HFREP_WIDE3_EPT") ? atoi(getenv("HFREP_WIDE3_EPT")) : 2;     // (1: measured equal — 28.3 vs 28.7 ms on 592 windows)
            const size_t smem = sizeof(float) * ((size_t)(2 + W3_NCH) * W3_BP * 8 + 2 * W3_BP + 32 + (size_t)W3_BP * W3_TS +
                                                 2 * (size_t)W3_EP * 8 + 64 * (size_t)256);
            if (ept == 2) {
                HFREP_CUDA_OK(allow_max_dynamic_smem(fit_wide3_kernel<2>));
                fit_wide3_kernel<2><<<n_tasks, 256, smem, st>>>(p);
            } else {
                HFREP_CUDA_OK(allow_max_dynamic_smem(fit_wide3_kernel<1>));
                fit_wide3_kernel<1><<<n_tasks, 512, smem, st>>>(p);
            }
            HFREP_CUDA_OK(cudaGetLastError());
            return HFREP_OK;
        }
        if (g.LP == 8) return launch_fit_wide2<8, 2>(p, g.NT, st);
        if (g.LP == 16) return launch_fit_wide2<16, 1>(p, g.NT, st);
        return launch_fit_wide2<32, 1>(p, g.NT, st);
    }
    const size_t smem = sizeof(float) * (2 * (size_t)c->batch * L + (size_t)c->batch * c->E + c->batch + 16);
    HFREP_REQUIRE(smem <= 227 * 1024, "fit_wide: %zu bytes of shared memory needed (batch x E too large)", smem);
    HFREP_CUDA_OK(allow_max_dynamic_smem(fit_wide_kernel));
    fit_wide_kernel<<<n_tasks, WIDE_THREADS, smem, (cudaStream_t)stream>>>(p);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_predict_wide(const double* etf, const double* hfd, const double* rf, int T, int E, int F, int L,
                                  int window, float alpha, int n_tasks, const int32_t* task_row0, const float* W,
                                  double* replica, int32_t* status, void* stream) {
    HFREP_REQUIRE(etf && hfd && rf && task_row0 && W && replica, "predict_wide: null argument");
    HFREP_REQUIRE(n_tasks >= 0, "predict_wide: n_tasks=%d", n_tasks);
    if (n_tasks == 0) return HFREP_OK;
    HFREP_REQUIRE(L >= 1 && L <= WIDE_MAX_L, "predict_wide: L=%d outside 1..%d", L, WIDE_MAX_L);
    HFREP_REQUIRE(window >= 2 && window <= 64, "predict_wide: window=%d outside 2..64", window);
    HFREP_REQUIRE(E >= 1 && F >= 1 && T >= window + 2, "predict_wide: E=%d F=%d T=%d", E, F, T);
    WideEvalParams p;
    p.etf = etf; p.hfd = hfd; p.rf = rf; p.task_row0 = task_row0; p.W = W; p.replica = replica; p.status = status;
    p.n_tasks = n_tasks; p.E = E; p.F = F; p.L = L; p.window = window; p.alpha = alpha;
    const size_t smem = sizeof(double) * (2 * (size_t)window * L + 3 * L) + sizeof(float) * (size_t)(window + 2) * L + E + 16;
    HFREP_CUDA_OK(allow_max_dynamic_smem(eval_wide_kernel));
    eval_wide_kernel<<<n_tasks, WIDE_THREADS, smem, (cudaStream_t)stream>>>(p);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}
