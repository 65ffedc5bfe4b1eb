// Kernel (b): batched autoencoder fit — one task (E -> L -> E bias-free LeakyReLU autoencoder,
// MSE, Nadam, validation tail, early stopping) per warp team, whole training run in one launch.
//
// Replaces Autoencoder (Autoencoder_encapsulate.py:19-35) and AE.train (:72-96) of the reference,
// i.e. the Keras fit loop; semantics restated in SURVEY.md App. A.1-A.6.
//
// Data layout (per team, shared memory, floats; every row start 16-byte aligned):
//   Xs  [(T+1)][XS=28]   scaled training block, columns >= E zero, row T all zero (padding rows point at it)
//   We  [EP=24][LP]      encoder kernel (in x out), LP = L rounded up to 4, padding zero (stays zero: its gradient is 0)
//   Wd  [LP][WDS=28]     decoder kernel
//   H   [BP][HS]         h = lrelu(x We) of the current mini-batch (position-major)
//   dZy [BP][XS]         dLoss/d(decoder pre-activation)
//   dZh [BP][HS]         dLoss/d(encoder pre-activation)
// Nadam's m and v never leave registers: each thread owns a fixed set of 4x4 weight tiles.
// Every contraction is a 4x4 register-tiled FP32 FFMA mini-GEMM fed by 128-bit shared loads
// (8 LDS.128 per 64 FFMA); strides 28 / LP+4 keep 8 consecutive rows on distinct bank groups.
#include "common.cuh"

#include <cmath>

namespace hfrep {

constexpr int EP = 24;    // padded ETF width
constexpr int XS = 28;    // row stride of Xs and dZy
constexpr int WDS = 28;   // row stride of Wd

struct FitParams {
    const float* X;
    const int* task_L;
    const int* task_x;
    const float* W0;
    const int* perm;
    const float4* sched;
    float* W_out;
    float* loss_hist;
    float* val_hist;
    int* epochs_run;
    int* status;
    float* grad_out;
    long long perm_task_stride;
    int n_tasks, E, T, n_fit, n_val, batch, BP, max_epochs, patience, early_stop, w_stride, shared_x, dump_step;
    float alpha, lr, b1, b2, eps;
};

template <int LP>
struct Lay {
    static constexpr int HS = (LP % 8 == 4) ? LP : LP + 4;
    static constexpr int CT = LP / 4;
    __host__ __device__ static size_t team_floats(int BP) {
        return (size_t)EP * LP + (size_t)LP * WDS + (size_t)BP * HS * 2 + (size_t)BP * XS + BP + 32;
    }
};
__host__ __device__ inline size_t x_floats(int T) { return (size_t)(T + 1) * XS; }

template <int WPT>
__device__ __forceinline__ void team_sync(int team) {
    if constexpr (WPT == 1) {
        __syncwarp();
    } else {
        asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "n"(WPT * 32) : "memory");
    }
}

template <int WPT>
__device__ __forceinline__ float team_sum(float v, float* red, int team, int tid) {
    v = warp_sum(v);
    if constexpr (WPT == 1) {
        return v;
    } else {
        if ((tid & 31) == 0) red[tid >> 5] = v;
        team_sync<WPT>(team);
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < WPT; ++w) s += red[w];
        team_sync<WPT>(team);
        return s;
    }
}

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// acc[i][j] += sum_kk a[i][kk] * b[kk][j]      (A rows along k, B rows along n)
__device__ __forceinline__ void mma_nn(float (&acc)[4][4], const float4 (&a)[4], const float4 (&b)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float av[4] = {a[i].x, a[i].y, a[i].z, a[i].w};
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            acc[i][0] = fmaf(av[kk], b[kk].x, acc[i][0]);
            acc[i][1] = fmaf(av[kk], b[kk].y, acc[i][1]);
            acc[i][2] = fmaf(av[kk], b[kk].z, acc[i][2]);
            acc[i][3] = fmaf(av[kk], b[kk].w, acc[i][3]);
        }
    }
}
// acc[i][j] += sum_kk a[i][kk] * b[j][kk]      (both operands along k)
__device__ __forceinline__ void mma_nt(float (&acc)[4][4], const float4 (&a)[4], const float4 (&b)[4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
            acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
            acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
            acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
        }
    }
}
// acc[i][j] += a[i] * b[j]                     (outer product, one k)
__device__ __forceinline__ void mma_tn(float (&acc)[4][4], const float4& a, const float4& b) {
    const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        acc[i][0] = fmaf(av[i], b.x, acc[i][0]);
        acc[i][1] = fmaf(av[i], b.y, acc[i][1]);
        acc[i][2] = fmaf(av[i], b.z, acc[i][2]);
        acc[i][3] = fmaf(av[i], b.w, acc[i][3]);
    }
}

template <int LP, int WPT, int TPC>
__global__ void __launch_bounds__(32 * WPT * TPC) fit_kernel(const FitParams p) {
    extern __shared__ __align__(16) float smem[];
    constexpr int NTH = 32 * WPT;
    constexpr int HS = Lay<LP>::HS;
    constexpr int CT = Lay<LP>::CT;
    constexpr int NT5 = 2 * 6 * CT;                    // weight tiles: dWd (CT x 6) then dWe (6 x CT)
    constexpr int MAXT5 = (NT5 + NTH - 1) / NTH;       // tiles owned per thread

    const int team = threadIdx.x / NTH;
    const int tid = threadIdx.x - team * NTH;
    const int task = blockIdx.x * TPC + team;
    const bool active = task < p.n_tasks;
    const int T = p.T, E = p.E, BP = p.BP;

    const size_t xf = x_floats(T);
    const size_t tf = Lay<LP>::team_floats(BP);
    float* Xs;
    float* base;
    if (p.shared_x) {
        Xs = smem;
        base = smem + xf + (size_t)team * tf;
    } else {
        Xs = smem + (size_t)team * (xf + tf);
        base = Xs + xf;
    }
    float* We = base;
    float* Wd = We + EP * LP;
    float* H = Wd + LP * WDS;
    float* dZh = H + BP * HS;
    float* dZy = dZh + BP * HS;
    int* rowidx = reinterpret_cast<int*>(dZy + BP * XS);
    float* red = reinterpret_cast<float*>(rowidx + BP);

    // ---- stage the training block (fused window load: rows of the scaled block -> padded tile)
    if (p.shared_x) {
        const int n = (T + 1) * XS;
        for (int idx = threadIdx.x; idx < n; idx += blockDim.x) {
            const int r = idx / XS, c = idx - r * XS;
            Xs[idx] = (r < T && c < E) ? __ldg(p.X + (size_t)r * E + c) : 0.f;
        }
        __syncthreads();
        if (!active) return;
    } else {
        if (!active) return;
        const int xb = p.task_x ? p.task_x[task] : 0;
        const float* xg = p.X + (size_t)xb * T * E;
        const int n = (T + 1) * XS;
        for (int idx = tid; idx < n; idx += NTH) {
            const int r = idx / XS, c = idx - r * XS;
            Xs[idx] = (r < T && c < E) ? __ldg(xg + (size_t)r * E + c) : 0.f;
        }
    }

    const int L = p.task_L[task];
    const float* w0 = p.W0 + (size_t)task * p.w_stride;
    for (int idx = tid; idx < EP * LP; idx += NTH) {
        const int e = idx / LP, l = idx - e * LP;
        We[idx] = (e < E && l < L) ? w0[e * L + l] : 0.f;
    }
    for (int idx = tid; idx < LP * WDS; idx += NTH) {
        const int l = idx / WDS, e = idx - l * WDS;
        Wd[idx] = (l < L && e < E) ? w0[E * L + l * E + e] : 0.f;
    }

    float mom[MAXT5][16], vel[MAXT5][16];
#pragma unroll
    for (int j = 0; j < MAXT5; ++j)
#pragma unroll
        for (int q = 0; q < 16; ++q) mom[j][q] = vel[j][q] = 0.f;

    const int* perm_t = p.perm ? p.perm + (size_t)task * p.perm_task_stride : nullptr;
    const float alpha = p.alpha;
    const float omb1 = 1.f - p.b1, omb2 = 1.f - p.b2;

    // ---------------------------------------------------------------- phases
    // P1: H = lrelu(X_b We)            [nb x 24] . [24 x LP]
    auto fwd_encode = [&](int RT) {
        const int ntiles = RT * CT;
        for (int t = tid; t < ntiles; t += NTH) {
            const int ct = t / RT, rt = t - ct * RT;
            const float* ar[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) ar[i] = Xs + rowidx[rt + RT * i] * XS;
            const float* bp = We + ct * 4;
            float acc[4][4] = {};
#pragma unroll
            for (int k4 = 0; k4 < EP / 4; ++k4) {
                float4 a[4], b[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = ld4(ar[i] + 4 * k4);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) b[kk] = ld4(bp + (4 * k4 + kk) * LP);
                mma_nn(acc, a, b);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float4 h;
                h.x = lrelu(acc[i][0], alpha);
                h.y = lrelu(acc[i][1], alpha);
                h.z = lrelu(acc[i][2], alpha);
                h.w = lrelu(acc[i][3], alpha);
                st4(H + (rt + RT * i) * HS + ct * 4, h);
            }
        }
    };
    // P2: Zy = H Wd, y = lrelu(Zy), squared error, dZy   [nb x LP] . [LP x 24]
    auto fwd_decode = [&](int RT, bool train, float coef) -> float {
        float sse = 0.f;
        const int ntiles = RT * (EP / 4);
        for (int t = tid; t < ntiles; t += NTH) {
            const int ct = t / RT, rt = t - ct * RT;
            const float* bp = Wd + ct * 4;
            float acc[4][4] = {};
#pragma unroll
            for (int k4 = 0; k4 < CT; ++k4) {
                float4 a[4], b[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = ld4(H + (rt + RT * i) * HS + 4 * k4);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) b[kk] = ld4(bp + (4 * k4 + kk) * WDS);
                mma_nn(acc, a, b);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int pos = rt + RT * i;
                const float4 x = ld4(Xs + rowidx[pos] * XS + ct * 4);
                const float xv[4] = {x.x, x.y, x.z, x.w};
                float g[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float zy = acc[i][j];
                    const float d = lrelu(zy, alpha) - xv[j];
                    sse = fmaf(d, d, sse);
                    g[j] = lrelu_grad(zy, coef * d, alpha);
                }
                if (train) st4(dZy + pos * XS + ct * 4, make_float4(g[0], g[1], g[2], g[3]));
            }
        }
        return sse;
    };
    // P4: dZh = (dZy Wd^T) o lrelu'(Zh)   [nb x 24] . [24 x LP]
    auto bwd_hidden = [&](int RT) {
        const int ntiles = RT * CT;
        for (int t = tid; t < ntiles; t += NTH) {
            const int ct = t / RT, rt = t - ct * RT;
            float acc[4][4] = {};
#pragma unroll
            for (int k4 = 0; k4 < EP / 4; ++k4) {
                float4 a[4], b[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = ld4(dZy + (rt + RT * i) * XS + 4 * k4);
#pragma unroll
                for (int j = 0; j < 4; ++j) b[j] = ld4(Wd + (ct * 4 + j) * WDS + 4 * k4);
                mma_nt(acc, a, b);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int pos = rt + RT * i;
                const float4 h = ld4(H + pos * HS + ct * 4);     // sign(h) == sign(zh) for alpha > 0
                float4 g;
                g.x = lrelu_grad(h.x, acc[i][0], alpha);
                g.y = lrelu_grad(h.y, acc[i][1], alpha);
                g.z = lrelu_grad(h.z, acc[i][2], alpha);
                g.w = lrelu_grad(h.w, acc[i][3], alpha);
                st4(dZh + pos * HS + ct * 4, g);
            }
        }
    };
    // P5: dWd = H^T dZy, dWe = X_b^T dZh, then the Nadam update of the owned tiles
    auto grads_and_update = [&](int nb, int step) {
        const float4 sc = __ldg(p.sched + step);
        const bool dump = p.grad_out != nullptr && step == p.dump_step;
#pragma unroll
        for (int j = 0; j < MAXT5; ++j) {
            const int t = tid + j * NTH;
            if (t < NT5) {
                float acc[4][4] = {};
                float* wt;
                int wstride, r0, c0;
                const bool dec = t < 6 * CT;
                if (dec) {                                   // dWd tile: rows l (lt), cols e (et)
                    const int lt = t / 6, et = t - lt * 6;
                    const float* ap = H + lt * 4;
                    const float* bp = dZy + et * 4;
#pragma unroll 4
                    for (int q = 0; q < nb; ++q) mma_tn(acc, ld4(ap + q * HS), ld4(bp + q * XS));
                    wt = Wd + (lt * 4) * WDS + et * 4;
                    wstride = WDS; r0 = lt * 4; c0 = et * 4;
                } else {                                     // dWe tile: rows e (et), cols l (lt)
                    const int u = t - 6 * CT;
                    const int et = u / CT, lt = u - et * CT;
                    const float* ap = Xs + et * 4;
                    const float* bp = dZh + lt * 4;
#pragma unroll 4
                    for (int q = 0; q < nb; ++q) mma_tn(acc, ld4(ap + rowidx[q] * XS), ld4(bp + q * HS));
                    wt = We + (et * 4) * LP + lt * 4;
                    wstride = LP; r0 = et * 4; c0 = lt * 4;
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float4 w4 = ld4(wt + i * wstride);
                    float wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        const float g = acc[i][jj];
                        float& m = mom[j][i * 4 + jj];
                        float& v = vel[j][i * 4 + jj];
                        m = p.b1 * m + omb1 * g;
                        v = p.b2 * v + omb2 * (g * g);
                        const float m_bar = sc.x * g + sc.y * m;
                        wv[jj] -= p.lr * m_bar / (sqrtf(v * sc.z) + p.eps);
                    }
                    st4(wt + i * wstride, make_float4(wv[0], wv[1], wv[2], wv[3]));
                }
                if (dump) {
                    float* go = p.grad_out + (size_t)task * p.w_stride;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int jj = 0; jj < 4; ++jj) {
                            const int r = r0 + i, c = c0 + jj;
                            if (dec) { if (r < L && c < E) go[E * L + r * E + c] = acc[i][jj]; }
                            else     { if (r < E && c < L) go[r * L + c] = acc[i][jj]; }
                        }
                }
            }
        }
    };

    // ---------------------------------------------------------------- training run
    const int n_fit = p.n_fit, n_val = p.n_val, batch = p.batch;
    int step = 0, wait = 0, epochs_run = 0, st = 0;
    float best = INFINITY;
    for (int epoch = 0; epoch < p.max_epochs; ++epoch) {
        float lsum = 0.f;
        for (int lo = 0; lo < n_fit; lo += batch) {
            const int nb = min(batch, n_fit - lo);
            const int RT = (nb + 3) >> 2;
            for (int q = tid; q < 4 * RT; q += NTH)
                rowidx[q] = q < nb ? (perm_t ? __ldg(perm_t + (size_t)epoch * n_fit + lo + q) : lo + q) : T;
            team_sync<WPT>(team);
            fwd_encode(RT);
            team_sync<WPT>(team);
            lsum += fwd_decode(RT, true, 2.f / (float)(nb * E));
            team_sync<WPT>(team);
            bwd_hidden(RT);
            team_sync<WPT>(team);
            grads_and_update(nb, step);
            ++step;
            team_sync<WPT>(team);
        }
        float vsum = 0.f;
        for (int lo = 0; lo < n_val; lo += batch) {
            const int nb = min(batch, n_val - lo);
            const int RT = (nb + 3) >> 2;
            for (int q = tid; q < 4 * RT; q += NTH) rowidx[q] = q < nb ? n_fit + lo + q : T;
            team_sync<WPT>(team);
            fwd_encode(RT);
            team_sync<WPT>(team);
            vsum += fwd_decode(RT, false, 0.f);
            team_sync<WPT>(team);
        }
        const float loss = team_sum<WPT>(lsum, red, team, tid) / (float)(E * n_fit);
        const float val = n_val > 0 ? team_sum<WPT>(vsum, red, team, tid) / (float)(E * n_val) : NAN;
        if (tid == 0) {
            p.loss_hist[(size_t)task * p.max_epochs + epoch] = loss;
            p.val_hist[(size_t)task * p.max_epochs + epoch] = val;
        }
        epochs_run = epoch + 1;
        if (!isfinite(loss)) st = 1;
        if (p.early_stop) {                      // EarlyStopping.on_epoch_end (App. A.6)
            ++wait;
            if (val < best) { best = val; wait = 0; }
            if (wait >= p.patience && epoch > 0) break;
        }
    }

    float* wo = p.W_out + (size_t)task * p.w_stride;
    for (int idx = tid; idx < E * L; idx += NTH) {
        const int e = idx / L, l = idx - e * L;
        wo[idx] = We[e * LP + l];
    }
    for (int idx = tid; idx < L * E; idx += NTH) {
        const int l = idx / E, e = idx - l * E;
        wo[E * L + idx] = Wd[l * WDS + e];
    }
    if (tid == 0) {
        p.epochs_run[task] = epochs_run;
        p.status[task] = st;
    }
}

// ---------------------------------------------------------------------------------- host side
template <int LP, int WPT, int TPC>
static int launch_fit(const FitParams& p, cudaStream_t stream) {
    const size_t xf = x_floats(p.T), tf = Lay<LP>::team_floats(p.BP);
    const size_t smem = sizeof(float) * (p.shared_x ? xf + TPC * tf : TPC * (xf + tf));
    HFREP_REQUIRE(smem <= 227 * 1024, "fit: %zu bytes of shared memory needed (T=%d too long for the narrow kernel)",
                  smem, p.T);
    auto kern = fit_kernel<LP, WPT, TPC>;
    HFREP_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int grid = (p.n_tasks + TPC - 1) / TPC;
    kern<<<grid, 32 * WPT * TPC, smem, stream>>>(p);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

template <int LP>
static int dispatch_mapping(const FitParams& p, int wpt, cudaStream_t stream) {
    if (wpt == 1) return p.shared_x ? launch_fit<LP, 1, 8>(p, stream) : launch_fit<LP, 1, 1>(p, stream);
    return launch_fit<LP, 4, 1>(p, stream);
}

int fit_dispatch(const FitParams& p, int L_max, int wpt, cudaStream_t stream) {
    const int LP = (L_max + 3) & ~3;
    switch (LP) {
        case 4: return dispatch_mapping<4>(p, wpt, stream);
        case 8: return dispatch_mapping<8>(p, wpt, stream);
        case 12: return dispatch_mapping<12>(p, wpt, stream);
        case 16: return dispatch_mapping<16>(p, wpt, stream);
        case 20: return dispatch_mapping<20>(p, wpt, stream);
        case 24: return dispatch_mapping<24>(p, wpt, stream);
    }
    return set_err(HFREP_ERR_ARG, "fit: L_max=%d outside 1..%d", L_max, HFREP_MAX_L);
}

}  // namespace hfrep

using namespace hfrep;

extern "C" void hfrep_fit_cfg_default(hfrep_fit_cfg* c) {
    // literals of Autoencoder_encapsulate.py:19-35, :79-96 and Keras 2.7 Nadam() defaults
    memset(c, 0, sizeof(*c));
    c->E = 22; c->L_max = 8; c->T = 168; c->n_fit = 126; c->batch = 48; c->max_epochs = 1000;
    c->patience = 5; c->early_stop = 1; c->alpha = .2f;
    c->lr = 1e-3f; c->beta_1 = .9f; c->beta_2 = .999f; c->epsilon = 1e-7f; c->schedule_decay = .004f; c->decay_base = .96f;
    c->n_x = 1; c->warps_per_task = 0; c->dump_step = 0;
}

extern "C" int hfrep_nadam_schedule(const hfrep_fit_cfg* c, int n_steps, float* out) {
    HFREP_REQUIRE(c && out && n_steps >= 0, "nadam_schedule: null argument");
    // Keras 2.7 Nadam._prepare_local in float32
    float m_cache = 1.f;
    for (int s = 0; s < n_steps; ++s) {
        const float t = (float)(s + 1), nxt = (float)(s + 2);
        // float32 pow taken as the correctly rounded value (double pow, then rounded)
        auto pw = [](float x, float y) { return (float)pow((double)x, (double)y); };
        const float mu_t = c->beta_1 * (1.f - .5f * pw(c->decay_base, c->schedule_decay * t));
        const float mu_t1 = c->beta_1 * (1.f - .5f * pw(c->decay_base, c->schedule_decay * nxt));
        const float ms_new = m_cache * mu_t;
        const float ms_next = ms_new * mu_t1;
        out[4 * s + 0] = (1.f - mu_t) / (1.f - ms_new);
        out[4 * s + 1] = mu_t1 / (1.f - ms_next);
        out[4 * s + 2] = 1.f / (1.f - pw(c->beta_2, t));
        out[4 * s + 3] = mu_t;
        m_cache = ms_new;
    }
    return HFREP_OK;
}

extern "C" int hfrep_fit_batched(const hfrep_fit_cfg* c, const float* X, int n_tasks, const int32_t* task_L,
                                 const int32_t* task_x, const float* W0, int w_stride, const int32_t* perm,
                                 int64_t perm_task_stride, const float* sched, float* W_out, float* loss_hist,
                                 float* val_hist, int32_t* epochs_run, int32_t* status, float* grad_out,
                                 void* stream) {
    HFREP_REQUIRE(c && X && task_L && W0 && sched && W_out && loss_hist && val_hist && epochs_run && status,
                  "fit_batched: null argument");
    HFREP_REQUIRE(n_tasks >= 0, "fit_batched: n_tasks=%d", n_tasks);
    if (n_tasks == 0) return HFREP_OK;
    HFREP_REQUIRE(c->E >= 1 && c->E <= HFREP_MAX_E, "fit_batched: E=%d outside 1..%d", c->E, HFREP_MAX_E);
    HFREP_REQUIRE(c->L_max >= 1 && c->L_max <= HFREP_MAX_L, "fit_batched: L_max=%d outside 1..%d", c->L_max, HFREP_MAX_L);
    HFREP_REQUIRE(c->T >= 2 && c->n_fit >= 1 && c->n_fit <= c->T, "fit_batched: T=%d n_fit=%d", c->T, c->n_fit);
    HFREP_REQUIRE(c->batch >= 1 && c->batch <= 256, "fit_batched: batch=%d outside 1..256", c->batch);
    HFREP_REQUIRE(c->max_epochs >= 1, "fit_batched: max_epochs=%d", c->max_epochs);
    HFREP_REQUIRE(w_stride >= 2 * c->E * c->L_max, "fit_batched: w_stride=%d < 2*E*L_max", w_stride);
    HFREP_REQUIRE(c->n_x >= 1 && (c->n_x == 1 || task_x), "fit_batched: n_x=%d needs task_x", c->n_x);
    HFREP_REQUIRE(((uintptr_t)sched & 15) == 0, "fit_batched: sched must be 16-byte aligned");

    FitParams p;
    p.X = X; p.task_L = task_L; p.task_x = task_x; p.W0 = W0; p.perm = perm;
    p.sched = reinterpret_cast<const float4*>(sched);
    p.W_out = W_out; p.loss_hist = loss_hist; p.val_hist = val_hist; p.epochs_run = epochs_run; p.status = status;
    p.grad_out = grad_out; p.perm_task_stride = perm_task_stride;
    p.n_tasks = n_tasks; p.E = c->E; p.T = c->T; p.n_fit = c->n_fit; p.n_val = c->T - c->n_fit; p.batch = c->batch;
    p.BP = (c->batch + 3) & ~3; p.max_epochs = c->max_epochs; p.patience = c->patience; p.early_stop = c->early_stop;
    p.w_stride = w_stride; p.shared_x = (c->n_x == 1); p.dump_step = c->dump_step;
    p.alpha = c->alpha; p.lr = c->lr; p.b1 = c->beta_1; p.b2 = c->beta_2; p.eps = c->epsilon;

    int wpt = c->warps_per_task;
    if (wpt == 0) wpt = (n_tasks >= 4 * sm_count()) ? 1 : 4;    // few tasks: spend 4 warps on each to cut latency
    HFREP_REQUIRE(wpt == 1 || wpt == 4, "fit_batched: warps_per_task=%d (0, 1 or 4)", wpt);
    return fit_dispatch(p, c->L_max, wpt, (cudaStream_t)stream);
}
