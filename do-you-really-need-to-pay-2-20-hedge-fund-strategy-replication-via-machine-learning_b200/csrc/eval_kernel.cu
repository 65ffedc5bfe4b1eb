// Kernel (c): fused out-of-sample predict + rolling-window OLS + replica + ex-post + metrics.
// One trained autoencoder (task) per CTA of 256 threads; everything between the trained kernels
// and the scored replica series happens in one launch, intermediate results in shared memory.
//
// Replaces, in this order (reference lines in Autoencoder_encapsulate.py unless noted):
//   quality   model_IS_r2/RMSE :107-113, model_OOS_r2/RMSE :115-131 (167 expanding MinMax prefixes)
//   latent    encoder.predict on the raw test block :140
//   ols       OLS(Y, X).fit().params for window 0 :148-156 + helper.normalization helper.py:10-17
//   weights   LeakyReLU mask + beta_0/norm_0 mapping :158-169, pop last :179-180
//   ante      replica returns :183-193
//   post      helper.ex_post_return / transaction_cost / price_impact helper.py:65-92,112-131
//   turnover  :210-224
//   metrics   annualized_sharpe_ratio autoencoder_v4.ipynb:715-721; tracking error / correlation (ours, App. A.9)
//
// Precision follows the reference: network forward passes in float32 (Keras), everything else
// float64 (numpy/pandas/statsmodels).  OLS = Householder QR in float64 (the reference's pinv-OLS is
// the same least-squares solution for full-rank windows; rank deficiency is reported in status).
#include "common.cuh"

#include <cmath>

namespace hfrep {

constexpr int NTH_EVAL = 256;
constexpr int WS = 24;            // stride of We / Wd / M rows in shared memory (floats)

struct EvalParams {
    const double* Xtr; const double* Xte; const double* Yte; const double* rf; const double* sigma;
    const int* task_L; const int* task_row0; const float* W;
    double* ante; double* post; double* turnover; double* metrics; double* quality;
    unsigned char* mask; double* weights; double* beta; double* norm; int* status;
    int n_tasks, w_stride, E, F, T_train, T_test, window, want_oos, n_oos, n_pre, acols;
    int rolling, nq, wsz; // beta_rolling: OLS of every window (nq warps, one workspace of wsz doubles each)
    int series_dbl;       // doubles reserved for ante/post/turnover [3][n_oos][F] (>= the quality phase's prediction buffers)
    int qsplit;           // CTAs per task: 1, or part 0 builds the replica while parts 1.. share the quality work
    int off_float;        // byte offset of the float32 region inside dynamic shared memory (16-byte aligned)
    float alpha;
    double cost_param, phi;
};

// float32 forward pass of one row held in registers: y = lrelu(lrelu(x We) Wd)   (:32-35, predict)
__device__ __forceinline__ void forward_row(const float (&x)[HFREP_MAX_E], float (&y)[HFREP_MAX_E],
                                            const float* We, const float* Wd, int E, int L, float alpha) {
    float h[HFREP_MAX_L];
#pragma unroll
    for (int l = 0; l < HFREP_MAX_L; ++l) h[l] = 0.f;
#pragma unroll
    for (int e = 0; e < HFREP_MAX_E; ++e) {
        if (e < E) {
            const float xe = x[e];
#pragma unroll
            for (int l4 = 0; l4 < HFREP_MAX_L / 4; ++l4) {
                if (l4 * 4 < L) {
                    const float4 w = *reinterpret_cast<const float4*>(We + e * WS + l4 * 4);
                    h[l4 * 4 + 0] = fmaf(xe, w.x, h[l4 * 4 + 0]);
                    h[l4 * 4 + 1] = fmaf(xe, w.y, h[l4 * 4 + 1]);
                    h[l4 * 4 + 2] = fmaf(xe, w.z, h[l4 * 4 + 2]);
                    h[l4 * 4 + 3] = fmaf(xe, w.w, h[l4 * 4 + 3]);
                }
            }
        }
    }
#pragma unroll
    for (int e = 0; e < HFREP_MAX_E; ++e) y[e] = 0.f;
#pragma unroll
    for (int l = 0; l < HFREP_MAX_L; ++l) {
        if (l < L) {
            const float hl = lrelu(h[l], alpha);
#pragma unroll
            for (int e4 = 0; e4 < HFREP_MAX_E / 4; ++e4) {
                const float4 w = *reinterpret_cast<const float4*>(Wd + l * WS + e4 * 4);
                y[e4 * 4 + 0] = fmaf(hl, w.x, y[e4 * 4 + 0]);
                y[e4 * 4 + 1] = fmaf(hl, w.y, y[e4 * 4 + 1]);
                y[e4 * 4 + 2] = fmaf(hl, w.z, y[e4 * 4 + 2]);
                y[e4 * 4 + 3] = fmaf(hl, w.w, y[e4 * 4 + 3]);
            }
        }
    }
#pragma unroll
    for (int e = 0; e < HFREP_MAX_E; ++e) y[e] = lrelu(y[e], alpha);
}

// r2_score (uniform average, sklearn's zero-denominator rule) and per-column-RMSE mean of one block of `n` rows by one
// warp.  Pass of 32 rows: lane = row does the float32 forward pass in registers and parks the prediction in the warp's
// shared buffer ybuf [32][WS]; then lane = column accumulates that column's squared residuals and squared deviations in
// float64 (scaled value recomputed as x*scale + shift, MinMaxScaler, without FMA contraction).  No per-lane arrays of
// partial sums, so the kernel stays under 128 registers.
// cstat: per-warp shared scratch [3][HFREP_MAX_E] doubles = scale, shift, mean of the scaled column.
__device__ void block_quality(const double* x, int n, int E, int L, const double* cstat, const float* We,
                              const float* Wd, float alpha, int lane, float* ybuf, double& r2, double& rmse) {
    double ssr = 0., sst = 0.;                       // lane = column e
    const double sc = lane < E ? cstat[lane] : 0., sh = lane < E ? cstat[HFREP_MAX_E + lane] : 0.;
    const double mu = lane < E ? cstat[2 * HFREP_MAX_E + lane] : 0.;
    for (int r0 = 0; r0 < n; r0 += 32) {
        const int r = r0 + lane;
        if (r < n) {
            float xf[HFREP_MAX_E], yf[HFREP_MAX_E];
#pragma unroll
            for (int e = 0; e < HFREP_MAX_E; ++e)
                xf[e] = e < E ? (float)__dadd_rn(__dmul_rn(x[(size_t)r * E + e], cstat[e]), cstat[HFREP_MAX_E + e]) : 0.f;
            forward_row(xf, yf, We, Wd, E, L, alpha);
#pragma unroll
            for (int e4 = 0; e4 < HFREP_MAX_E / 4; ++e4)
                *reinterpret_cast<float4*>(ybuf + lane * WS + 4 * e4) = make_float4(yf[4 * e4], yf[4 * e4 + 1], yf[4 * e4 + 2], yf[4 * e4 + 3]);
        }
        __syncwarp();
        if (lane < E) {
            const int m = min(32, n - r0);
            for (int q = 0; q < m; ++q) {
                const double xs = __dadd_rn(__dmul_rn(x[(size_t)(r0 + q) * E + lane], sc), sh);   // X *= scale; X += min_
                const double d = xs - (double)ybuf[q * WS + lane];
                const double c = xs - mu;
                ssr += d * d;
                sst += c * c;
            }
        }
        __syncwarp();
    }
    const double r2e = lane < E ? (sst != 0. ? 1. - ssr / sst : (ssr != 0. ? 0. : 1.)) : 0.;
    const double rme = lane < E ? sqrt(ssr / n) : 0.;
    r2 = warp_sum(r2e) / E;
    rmse = warp_sum(rme) / E;
}

// The 11 per-series statistics of one return series of n values (stride ss) against a fund series y (stride ys),
// computed by one warp; lane 0 writes out[j*ostride], j = 0..HFREP_N_STATS-1:
//   0 Sharpe (autoencoder_v4.ipynb:715-721: (mean r - mean rf)/std(r, ddof=0)*sqrt(12))
//   1 tracking error, 2 correlation (not in the reference)
//   3 Omega(0 %), 4 Omega(10 %)  (ipynb:695-705, threshold de-annualised as (1+thr)^sqrt(1/252)-1)
//   5 historical cVaR(5 %)       (ipynb:729-757: mean of r <= np.percentile(r, 5), linear interpolation)
//   6..8 CEQ(gamma = 2, 5, 10)   (ipynb:759-770)
//   9 skewness, 10 excess kurtosis (pandas DataFrame.skew()/kurt(), bias-corrected; ipynb:836)
__device__ void series_stats(const double* s, int ss, const double* y, int ys, const double* rf, int n, int lane,
                             double* out, int ostride) {
    double sr = 0., sy = 0., sf = 0.;
    for (int k = lane; k < n; k += 32) { sr += s[(size_t)k * ss]; sy += y[(size_t)k * ys]; sf += rf[k]; }
    const double mr = warp_sum(sr) / n, my = warp_sum(sy) / n, rfm = warp_sum(sf) / n;
    const double thr0 = pow(1., sqrt(1. / 252.)) - 1., thr10 = pow(1.1, sqrt(1. / 252.)) - 1.;
    double vrr = 0., vyy = 0., vry = 0., sd = 0., m3 = 0., m4 = 0.;
    double op0 = 0., on0 = 0., op10 = 0., on10 = 0., c2 = 0., c5 = 0., c10 = 0.;
    for (int k = lane; k < n; k += 32) {
        const double r = s[(size_t)k * ss];
        const double a = r - mr, b = y[(size_t)k * ys] - my;
        vrr += a * a; vyy += b * b; vry += a * b; sd += (a - b) * (a - b);
        m3 += a * a * a; m4 += a * a * a * a;
        const double e0 = r - thr0, e10 = r - thr10;
        if (e0 > 0.) op0 += e0; else if (e0 < 0.) on0 += e0;
        if (e10 > 0.) op10 += e10; else if (e10 < 0.) on10 += e10;
        const double g = (1. + r) / (1. + rf[k]);
        c2 += pow(g, -1.); c5 += pow(g, -4.); c10 += pow(g, -9.);
    }
    vrr = warp_sum(vrr); vyy = warp_sum(vyy); vry = warp_sum(vry); sd = warp_sum(sd);
    m3 = warp_sum(m3); m4 = warp_sum(m4);
    op0 = warp_sum(op0); on0 = warp_sum(on0); op10 = warp_sum(op10); on10 = warp_sum(on10);
    c2 = warp_sum(c2); c5 = warp_sum(c5); c10 = warp_sum(c10);
    // np.percentile(r, 5): order statistics lo, lo+1 by (stable) rank counting, then linear interpolation
    const double pos = (n - 1) * 0.05;
    const int lo = (int)floor(pos);
    const double t = pos - lo;
    double slo = 0., shi = 0.;
    for (int i = lane; i < n; i += 32) {
        const double v = s[(size_t)i * ss];
        int rank = 0;
        for (int j = 0; j < n; ++j) {
            const double u = s[(size_t)j * ss];
            rank += (u < v) || (u == v && j < i);
        }
        if (rank == lo) slo = v;
        if (rank == min(lo + 1, n - 1)) shi = v;
    }
    slo = warp_sum(slo); shi = warp_sum(shi);          // exactly one lane holds each
    const double d = shi - slo;
    const double q = t >= .5 ? shi - d * (1. - t) : slo + d * t;     // numpy's _lerp
    double cs = 0., cn = 0.;
    for (int k = lane; k < n; k += 32) {
        const double r = s[(size_t)k * ss];
        if (r <= q) { cs += r; cn += 1.; }
    }
    cs = warp_sum(cs); cn = warp_sum(cn);
    if (lane == 0) {
        const double dn = (double)n;
        const double m2 = vrr / dn;
        out[0 * ostride] = (mr - rfm) / sqrt(m2) * sqrt(12.);
        out[1 * ostride] = sqrt(sd / (dn - 1.)) * sqrt(12.);
        out[2 * ostride] = vry / sqrt(vrr * vyy);
        out[3 * ostride] = op0 / (-on0);
        out[4 * ostride] = op10 / (-on10);
        out[5 * ostride] = cs / cn;
        out[6 * ostride] = log(c2 / dn) / ((1. - 2.) / 12.);
        out[7 * ostride] = log(c5 / dn) / ((1. - 5.) / 12.);
        out[8 * ostride] = log(c10 / dn) / ((1. - 10.) / 12.);
        out[9 * ostride] = sqrt(dn * (dn - 1.)) / (dn - 2.) * (m3 / dn) / pow(m2, 1.5);
        const double g2 = (m4 / dn) / (m2 * m2) - 3.;
        out[10 * ostride] = (dn - 1.) / ((dn - 2.) * (dn - 3.)) * ((dn + 1.) * g2 + 6.);
    }
}

__global__ void __launch_bounds__(NTH_EVAL, 2) eval_kernel(const EvalParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int task = blockIdx.x / p.qsplit, qpart = blockIdx.x - task * p.qsplit;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = NTH_EVAL / 32;
    const int E = p.E, F = p.F, L = p.task_L[task], window = p.window, n_oos = p.n_oos, T_test = p.T_test;
    const float alpha = p.alpha;
    // rolling refit: every task evaluates its own T_test rows of the full matrices, starting at task_row0[task]
    const int row0 = p.task_row0 ? p.task_row0[task] : 0;
    const double* Xte = p.Xte + (size_t)row0 * E;
    const double* Yte = p.Yte + (size_t)row0 * F;
    const double* rfv = p.rf + row0;

    // ---- shared memory carve-up (doubles first)
    // OLS workspaces: nq blocks of wsz doubles = [A (window x acols) | beta (MAX_L x F) | G (F x E) | norm (F) | vhh (window) |
    //                                             scratch of the rank-deficient path (MAX_L x MAX_L + 3 MAX_L)]
    double* wsp = reinterpret_cast<double*>(smem_raw);
    double* betaS = wsp + (size_t)window * p.acols;               // workspace 0: beta_0, G_0, norm_0 (the reference's choice)
    double* G = betaS + HFREP_MAX_L * F;
    double* normS = G + F * E;
    double* anteS = wsp + (size_t)p.nq * p.wsz;                   // [n_oos][F]
    double* postS = anteS + (size_t)n_oos * F;                    // [n_oos][F]
    double* turnS = postS + (size_t)n_oos * F;                    // [n_oos][F]
    double* cstat = anteS + p.series_dbl;                         // [NW][3][HFREP_MAX_E]
    // the quality phase runs before the replica stages: its prediction buffers [NW][32][WS] floats share their storage
    float* ybuf = reinterpret_cast<float*>(anteS) + warp * 32 * WS;
    float* We = reinterpret_cast<float*>(smem_raw + p.off_float); // [E][WS]
    float* Wd = We + HFREP_MAX_E * WS;                            // [L][WS]
    float* M = Wd + HFREP_MAX_L * WS;                             // [T_test][WS] latent factors
    unsigned char* maskS = reinterpret_cast<unsigned char*>(M + (size_t)T_test * WS);   // [n_oos][E]
    __shared__ int s_status;

    const float* w = p.W + (size_t)task * p.w_stride;
    for (int idx = tid; idx < HFREP_MAX_E * WS; idx += NTH_EVAL) {
        const int e = idx / WS, l = idx - e * WS;
        We[idx] = (e < E && l < L) ? w[e * L + l] : 0.f;
    }
    for (int idx = tid; idx < HFREP_MAX_L * WS; idx += NTH_EVAL) {
        const int l = idx / WS, e = idx - l * WS;
        Wd[idx] = (l < L && e < E) ? w[E * L + l * E + e] : 0.f;
    }
    if (tid == 0) s_status = 0;
    __syncthreads();

    // ---- quality: in-sample block (warp 0) and the expanding prefixes (all warps, incremental column stats).  With
    // several CTAs per task the quality work belongs to CTAs 1.. and CTA 0 goes straight to the replica stages, so a
    // task's two halves run side by side (the latency of a handful of tasks is their sum otherwise).
    const int qq = p.qsplit > 1 ? qpart - 1 : 0, nqq = p.qsplit > 1 ? p.qsplit - 1 : 1;
    if (p.quality && (p.qsplit == 1 || qpart > 0)) {
        double* q = p.quality + (size_t)task * (2 + 2 * p.n_pre);
        double* cs = cstat + warp * 3 * HFREP_MAX_E;
        if (warp == 0 && qq == 0) {
            if (lane < HFREP_MAX_E) {
                double s = 0.;
                if (lane < E) for (int r = 0; r < p.T_train; ++r) s += p.Xtr[(size_t)r * E + lane];
                cs[lane] = 1.; cs[HFREP_MAX_E + lane] = 0.; cs[2 * HFREP_MAX_E + lane] = s / p.T_train;
            }
            __syncwarp();
            double r2, rmse;
            block_quality(p.Xtr, p.T_train, E, L, cs, We, Wd, alpha, lane, ybuf, r2, rmse);
            if (lane == 0) { q[0] = r2; q[1] = rmse; }
            __syncwarp();
        }
        if (p.want_oos) {
            double lo = INFINITY, hi = -INFINITY, sum = 0.;
            int seen = 0;                                         // rows [0, seen) folded into lo/hi/sum (lane = column)
            for (int j = qq * NW + warp; j < p.n_pre; j += NW * nqq) {
                const int n = j + 2;                              // prefix x_test[:n]
                if (lane < E)
                    for (; seen < n; ++seen) {
                        const double v = Xte[(size_t)seen * E + lane];
                        lo = fmin(lo, v); hi = fmax(hi, v); sum += v;
                    }
                seen = n;
                if (lane < HFREP_MAX_E) {
                    double scale = 1., shift = 0., mean = 0.;
                    if (lane < E) {
                        double rng = hi - lo;
                        if (rng == 0.) rng = 1.;
                        scale = 1. / rng;                         // MinMaxScaler: scale_ = 1/range, min_ = -min*scale_
                        shift = __dsub_rn(0., __dmul_rn(lo, scale));
                        mean = (sum / n) * scale + shift;
                    }
                    cs[lane] = scale; cs[HFREP_MAX_E + lane] = shift; cs[2 * HFREP_MAX_E + lane] = mean;
                }
                __syncwarp();
                double r2, rmse;
                block_quality(Xte, n, E, L, cs, We, Wd, alpha, lane, ybuf, r2, rmse);
                if (lane == 0) { q[2 + j] = r2; q[2 + p.n_pre + j] = rmse; }
                __syncwarp();
            }
        }
    }

    if (qpart != 0) return;          // the other CTAs of this task only score expanding prefixes

    // ---- latent factors of the raw test block, float32 (:140)
    for (int idx = tid; idx < T_test * WS; idx += NTH_EVAL) {
        const int r = idx / WS, l = idx - r * WS;
        float acc = 0.f;
        if (l < L)
            for (int e = 0; e < E; ++e) acc = fmaf((float)Xte[(size_t)r * E + e], We[e * WS + l], acc);
        M[idx] = lrelu(acc, alpha);
    }
    __syncthreads();

    // ---- LeakyReLU masks of the decoder pre-activation, float32 (:163-166), months 0..n_oos-1
    for (int idx = tid; idx < n_oos * E; idx += NTH_EVAL) {
        const int k = idx / E, e = idx - k * E;
        float acc = 0.f;
        for (int l = 0; l < L; ++l) acc = fmaf(M[(window + k) * WS + l], Wd[l * WS + e], acc);
        maskS[idx] = acc < 0.f ? 1 : 0;
    }

    // ---- OLS (:148-156): Householder QR of [M_w | Y_w] in float64 by one warp (lane = column), then
    //      helper.normalization (helper.py:10-17) and G_w = beta_w^T Wd (:167).  The reference only ever uses window 0
    //      (App. B.1); beta_rolling solves every window and writes that month's weights straight to `weights`.
    const int ncol = L + F;
    auto ols_window = [&](int w0, double* ws) {
        double* A = ws;
        double* beta = A + (size_t)window * p.acols;
        double* Gw = beta + HFREP_MAX_L * F;
        double* normW = Gw + F * E;
        double* vhh = normW + F;
        const float* Mw = M + (size_t)w0 * WS;
        const double* Yw = Yte + (size_t)w0 * F;
        for (int idx = lane; idx < window * ncol; idx += 32) {
            const int r = idx / ncol, c = idx - r * ncol;
            A[r * p.acols + c] = c < L ? (double)Mw[r * WS + c] : Yw[(size_t)r * F + (c - L)];
        }
        __syncwarp();
        double dmax = 0., dmin = INFINITY;
        for (int k = 0; k < L; ++k) {
            // column k below the diagonal -> Householder vector v (all lanes compute the same scalars)
            double sq = 0.;
            for (int r = k; r < window; ++r) { const double a = A[r * p.acols + k]; sq += a * a; }
            const double akk = A[k * p.acols + k];
            const double nrm = sqrt(sq);
            const double alph = akk > 0. ? -nrm : nrm;
            const double vtv = sq - 2. * akk * alph + alph * alph;      // |x - alph e1|^2
            dmax = fmax(dmax, fabs(alph)); dmin = fmin(dmin, fabs(alph));
            __syncwarp();
            for (int r = k + lane; r < window; r += 32) vhh[r] = A[r * p.acols + k] - (r == k ? alph : 0.);
            __syncwarp();
            if (vtv > 0.) {
                const double bh = 2. / vtv;
                for (int c = k + 1 + lane; c < ncol; c += 32) {
                    double wj = 0.;
                    for (int r = k; r < window; ++r) wj += vhh[r] * A[r * p.acols + c];
                    wj *= bh;
                    for (int r = k; r < window; ++r) A[r * p.acols + c] -= vhh[r] * wj;
                }
            }
            __syncwarp();
            if (lane == 0) {
                A[k * p.acols + k] = alph;
                for (int r = k + 1; r < window; ++r) A[r * p.acols + k] = 0.;
            }
            __syncwarp();
        }
        const bool deficient = !(dmin > 1e-13 * dmax);
        if (!deficient) {
            // back substitution, lane = fund
            for (int f = lane; f < F; f += 32)
                for (int l = L - 1; l >= 0; --l) {
                    double acc = A[l * p.acols + L + f];
                    for (int j = l + 1; j < L; ++j) acc -= A[l * p.acols + j] * beta[j * F + f];
                    beta[l * F + f] = acc / A[l * p.acols + l];
                }
        } else {
            // Rank-deficient latent block (a dead or duplicated latent unit): OLS(Y, X).fit() (:151) is pinv(X) Y, the
            // MINIMUM-NORM least-squares solution.  Complete orthogonal decomposition, as LAPACK's xGELSY does it:
            // column-pivoted Householder QR  M P = Q [T; 0]  with T = [R11 R12] of rank r rows, then the QR of T^T
            // (L x r) = Z [S; 0], so that  beta = P Z [S^-T (Q^T y)_1..r ; 0].  status = 2 reports that this path ran.
            if (lane == 0) s_status = 2;
            double* S2 = normW + F + window;              // [L][MAX_L]: T^T, then Householder vectors below S's diagonal
            double* d2 = S2 + HFREP_MAX_L * HFREP_MAX_L;  // [L] diagonal of S
            double* t2 = d2 + HFREP_MAX_L;                // [L] 2 / v^T v
            int* perm = reinterpret_cast<int*>(t2 + HFREP_MAX_L);   // [L] column order
            __syncwarp();
            for (int idx = lane; idx < window * ncol; idx += 32) {
                const int r = idx / ncol, c = idx - r * ncol;
                A[r * p.acols + c] = c < L ? (double)Mw[r * WS + c] : Yw[(size_t)r * F + (c - L)];
            }
            if (lane < L) perm[lane] = lane;
            __syncwarp();
            int rank = 0;
            double n0 = 0.;
            for (int k = 0; k < L; ++k) {
                // pivot: the remaining column with the largest norm below row k (every lane computes the same choice)
                int best = k;
                double bsq = -1.;
                for (int c = k; c < L; ++c) {
                    double sq = 0.;
                    for (int r = k; r < window; ++r) { const double a = A[r * p.acols + c]; sq += a * a; }
                    if (sq > bsq) { bsq = sq; best = c; }
                }
                const double nrm = sqrt(bsq);
                if (k == 0) n0 = nrm;
                if (!(nrm > 1e-13 * n0)) break;           // what is left is numerically zero: rank = k
                __syncwarp();
                if (best != k) {
                    for (int r = lane; r < window; r += 32) {
                        const double t = A[r * p.acols + k]; A[r * p.acols + k] = A[r * p.acols + best]; A[r * p.acols + best] = t;
                    }
                    if (lane == 0) { const int t = perm[k]; perm[k] = perm[best]; perm[best] = t; }
                }
                __syncwarp();
                const double akk = A[k * p.acols + k];
                const double alph = akk > 0. ? -nrm : nrm;
                const double vtv = bsq - 2. * akk * alph + alph * alph;
                __syncwarp();
                for (int r = k + lane; r < window; r += 32) vhh[r] = A[r * p.acols + k] - (r == k ? alph : 0.);
                __syncwarp();
                const double bh = 2. / vtv;
                for (int c = k + 1 + lane; c < ncol; c += 32) {
                    double wj = 0.;
                    for (int r = k; r < window; ++r) wj += vhh[r] * A[r * p.acols + c];
                    wj *= bh;
                    for (int r = k; r < window; ++r) A[r * p.acols + c] -= vhh[r] * wj;
                }
                __syncwarp();
                if (lane == 0) {
                    A[k * p.acols + k] = alph;
                    for (int r = k + 1; r < window; ++r) A[r * p.acols + k] = 0.;
                }
                __syncwarp();
                rank = k + 1;
            }
            // T^T (L x rank) and its Householder QR
            for (int idx = lane; idx < L * rank; idx += 32) {
                const int i = idx / rank, j = idx - i * rank;               // S2[i][j] = T[j][i] (zero below T's diagonal)
                S2[i * HFREP_MAX_L + j] = i >= j ? A[j * p.acols + i] : 0.;
            }
            __syncwarp();
            for (int k = 0; k < rank; ++k) {
                double sq = 0.;
                for (int i = k; i < L; ++i) { const double a = S2[i * HFREP_MAX_L + k]; sq += a * a; }
                const double akk = S2[k * HFREP_MAX_L + k];
                const double nrm = sqrt(sq);
                const double alph = akk > 0. ? -nrm : nrm;
                const double vtv = sq - 2. * akk * alph + alph * alph;
                const double bh = vtv > 0. ? 2. / vtv : 0.;
                __syncwarp();
                if (lane == 0) { S2[k * HFREP_MAX_L + k] = akk - alph; d2[k] = alph; t2[k] = bh; }   // v = x - alph e1, in place
                __syncwarp();
                for (int c = k + 1 + lane; c < rank; c += 32) {
                    double wj = 0.;
                    for (int i = k; i < L; ++i) wj += S2[i * HFREP_MAX_L + k] * S2[i * HFREP_MAX_L + c];
                    wj *= bh;
                    for (int i = k; i < L; ++i) S2[i * HFREP_MAX_L + c] -= S2[i * HFREP_MAX_L + k] * wj;
                }
                __syncwarp();
            }
            for (int f = lane; f < F; f += 32) {
                double z[HFREP_MAX_L];
                for (int i = 0; i < rank; ++i) {                            // S^T u = (Q^T y)_1..r : forward substitution
                    double acc = A[i * p.acols + L + f];
                    for (int j = 0; j < i; ++j) acc -= S2[j * HFREP_MAX_L + i] * z[j];
                    z[i] = acc / d2[i];
                }
                for (int i = rank; i < L; ++i) z[i] = 0.;
                for (int k = rank - 1; k >= 0; --k) {                       // beta' = H_0 .. H_{r-1} [u; 0]
                    double wj = 0.;
                    for (int i = k; i < L; ++i) wj += S2[i * HFREP_MAX_L + k] * z[i];
                    wj *= t2[k];
                    for (int i = k; i < L; ++i) z[i] -= S2[i * HFREP_MAX_L + k] * wj;
                }
                for (int i = 0; i < L; ++i) beta[perm[i] * F + f] = z[i];   // undo the column order
            }
        }
        __syncwarp();
        for (int f = lane; f < F; f += 32) {
            // helper.normalization: sd(Y)/sd(M beta), ddof = 1
            double my = 0., mr = 0.;
            for (int r = 0; r < window; ++r) {
                double rh = 0.;
                for (int l = 0; l < L; ++l) rh += (double)Mw[r * WS + l] * beta[l * F + f];
                mr += rh; my += Yw[(size_t)r * F + f];
            }
            mr /= window; my /= window;
            double vr = 0., vy = 0.;
            for (int r = 0; r < window; ++r) {
                double rh = 0.;
                for (int l = 0; l < L; ++l) rh += (double)Mw[r * WS + l] * beta[l * F + f];
                const double dy = Yw[(size_t)r * F + f] - my;
                vr += (rh - mr) * (rh - mr) / (window - 1);
                vy += dy * dy / (window - 1);
            }
            normW[f] = sqrt(vy) / sqrt(vr);
        }
        __syncwarp();
        for (int idx = lane; idx < F * E; idx += 32) {            // G_w = beta_w^T Wd
            const int f = idx / E, e = idx - f * E;
            double acc = 0.;
            for (int l = 0; l < L; ++l) acc += beta[l * F + f] * (double)Wd[l * WS + e];
            Gw[idx] = acc;
        }
        __syncwarp();
    };
    double* wg = p.weights ? p.weights + (size_t)task * n_oos * E * F : nullptr;
    __syncthreads();                                              // masks are complete
    if (!p.rolling) {
        if (warp == 0) ols_window(0, wsp);
    } else if (warp < p.nq) {
        double* ws = wsp + (size_t)warp * p.wsz;
        for (int w = warp; w < n_oos; w += p.nq) {
            ols_window(w, ws);
            const double* Gw = ws + (size_t)window * p.acols + HFREP_MAX_L * F;
            const double* normW = Gw + F * E;
            for (int idx = lane; idx < E * F; idx += 32) {        // that month's weights, layout [e][f]
                const int e = idx / F, f = idx - e * F;
                wg[((size_t)w * E + e) * F + f] = Gw[f * E + e] * (maskS[w * E + e] ? .2 : 1.) * normW[f];
            }
            __syncwarp();
        }
        if (warp == 0) ols_window(0, wsp);                        // the reported beta / norm stay window 0's
    }
    __syncthreads();

    // ---- per (month, fund): weights, delta, ante (:167-168, :188-190), ex-post penalty (helper.py:120-129), |dw|
    const double scale_c = p.cost_param, phi = p.phi;
    for (int idx = tid; idx < n_oos * F; idx += NTH_EVAL) {
        const int k = idx / F, f = idx - k * F;
        const double nf = normS[f];
        const int row = window + 1 + k;                           // test-block row of OOS month k (App. B.3)
        double sumw = 0., dot = 0., pen = 0., dabs = 0.;
        for (int e = 0; e < E; ++e) {
            const double g = G[f * E + e];
            auto wt = [&](int kk) -> double {                     // weight of ETF e for fund f in month kk
                return p.rolling ? wg[((size_t)kk * E + e) * F + f] : g * (maskS[kk * E + e] ? .2 : 1.) * nf;
            };
            const double wk = wt(k);
            sumw += wk;
            dot += Xte[(size_t)row * E + e] * wk;
            if (wg && !p.rolling) wg[((size_t)k * E + e) * F + f] = wk;
            if (k >= 1 && p.sigma) {                              // transaction_cost + price_impact, helper.py:65-92
                const double old = wt(k - 1);
                const double sg = p.sigma[(size_t)(k - 1) * E + e] * scale_c;
                const double d = old - wk;
                const double tc = .5 * d * d * sg;
                const double pi = phi * wk * sg * d - old * sg * d - .5 * d * d * sg;
                pen += tc + pi;
            }
            if (k + 1 < n_oos) dabs += fabs(wk - wt(k + 1));
        }
        const double a = (1. - sumw) * rfv[row] + dot;
        anteS[idx] = a;
        postS[idx] = a + pen;                                     // post[0] = ante[0] (pen = 0)
        turnS[idx] = dabs;
    }
    __syncthreads();

    // ---- stores + per-fund reductions (warp per fund)
    if (p.ante) for (int idx = tid; idx < n_oos * F; idx += NTH_EVAL) p.ante[(size_t)task * n_oos * F + idx] = anteS[idx];
    if (p.post) for (int idx = tid; idx < n_oos * F; idx += NTH_EVAL) p.post[(size_t)task * n_oos * F + idx] = postS[idx];
    if (p.mask) for (int idx = tid; idx < n_oos * E; idx += NTH_EVAL) p.mask[(size_t)task * n_oos * E + idx] = maskS[idx];
    if (p.beta) for (int idx = tid; idx < HFREP_MAX_L * F; idx += NTH_EVAL)
        p.beta[(size_t)task * HFREP_MAX_L * F + idx] = idx < L * F ? betaS[idx] : 0.;
    if (p.norm) for (int idx = tid; idx < F; idx += NTH_EVAL) p.norm[(size_t)task * F + idx] = normS[idx];

    for (int f = warp; f < F; f += NW) {
        double ts = 0.;
        for (int k = lane; k < n_oos; k += 32) ts += turnS[k * F + f];
        ts = warp_sum(ts);
        if (lane == 0 && p.turnover) p.turnover[(size_t)task * F + f] = ts / (n_oos / 12.);     // :216
        if (p.metrics && n_oos >= 4) {
#pragma unroll 1
            for (int which = 0; which < 2; ++which)
                series_stats((which ? postS : anteS) + f, F, Yte + (size_t)(window + 1) * F + f, F, rfv + window + 1, n_oos,
                             lane, p.metrics + ((size_t)task * 2 + which) * HFREP_N_STATS * F + f, F);
        }
    }
    if (tid == 0 && p.status) p.status[task] = s_status;
}

// Forward pass of arbitrary rows, one thread per row; weights staged in shared memory.
__global__ void __launch_bounds__(128) forward_kernel(const float* __restrict__ W, int E, int L, const float* __restrict__ x,
                                                      int n, int encode_only, float alpha, float* __restrict__ out) {
    __shared__ __align__(16) float We[HFREP_MAX_E * WS];
    __shared__ __align__(16) float Wd[HFREP_MAX_L * WS];
    for (int idx = threadIdx.x; idx < HFREP_MAX_E * WS; idx += blockDim.x) {
        const int e = idx / WS, l = idx - e * WS;
        We[idx] = (e < E && l < L) ? W[e * L + l] : 0.f;
    }
    for (int idx = threadIdx.x; idx < HFREP_MAX_L * WS; idx += blockDim.x) {
        const int l = idx / WS, e = idx - l * WS;
        Wd[idx] = (l < L && e < E) ? W[E * L + l * E + e] : 0.f;
    }
    __syncthreads();
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    if (encode_only) {
        for (int l = 0; l < L; ++l) {
            float acc = 0.f;
            for (int e = 0; e < E; ++e) acc = fmaf(x[(size_t)r * E + e], We[e * WS + l], acc);
            out[(size_t)r * L + l] = lrelu(acc, alpha);
        }
    } else {
        float xf[HFREP_MAX_E], yf[HFREP_MAX_E];
#pragma unroll
        for (int e = 0; e < HFREP_MAX_E; ++e) xf[e] = e < E ? x[(size_t)r * E + e] : 0.f;
        forward_row(xf, yf, We, Wd, E, L, alpha);
#pragma unroll
        for (int e = 0; e < HFREP_MAX_E; ++e)
            if (e < E) out[(size_t)r * E + e] = yf[e];
    }
}

// Stand-alone series statistics: one warp per fund.
__global__ void series_metrics_kernel(const double* s, const double* y, const double* rf, int n, int F, double* out) {
    const int lane = threadIdx.x & 31;
    const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (f >= F) return;
    series_stats(s + f, F, y + f, F, rf, n, lane, out + f, F);
}

}  // namespace hfrep

using namespace hfrep;

extern "C" void hfrep_eval_cfg_default(hfrep_eval_cfg* c) {
    memset(c, 0, sizeof(*c));
    c->E = 22; c->F = 13; c->T_train = 168; c->T_test = 169; c->window = 24;     // autoencoder_v4.ipynb:112, :133
    c->beta_rolling = 0; c->want_oos = 1; c->alpha = .2f; c->cost_param = .05; c->phi = .5;   // helper.py:65,83
}

extern "C" int hfrep_predict_metrics(const hfrep_eval_cfg* c, const double* Xtr, const double* Xte, const double* Yte,
                                     const double* rf, const double* sigma, int n_tasks, const int32_t* task_L,
                                     const int32_t* task_row0, const float* W, int w_stride, double* ante, double* post, double* turnover,
                                     double* metrics, double* quality, uint8_t* mask, double* weights, double* beta,
                                     double* norm, int32_t* status, void* stream) {
    HFREP_REQUIRE(c && Xte && Yte && rf && task_L && W, "predict_metrics: null argument");
    HFREP_REQUIRE(n_tasks >= 0, "predict_metrics: n_tasks=%d", n_tasks);
    if (n_tasks == 0) return HFREP_OK;
    HFREP_REQUIRE(c->E >= 1 && c->E <= HFREP_MAX_E, "predict_metrics: E=%d outside 1..%d", c->E, HFREP_MAX_E);
    HFREP_REQUIRE(c->F >= 1 && c->F <= HFREP_MAX_F, "predict_metrics: F=%d outside 1..%d", c->F, HFREP_MAX_F);
    HFREP_REQUIRE(c->window >= 2 && c->window <= 64, "predict_metrics: window=%d outside 2..64", c->window);
    HFREP_REQUIRE(c->T_test >= c->window + 2, "predict_metrics: T_test=%d needs at least window+2 rows", c->T_test);
    HFREP_REQUIRE(w_stride >= 2 * c->E, "predict_metrics: w_stride=%d", w_stride);
    HFREP_REQUIRE(!c->beta_rolling || weights, "predict_metrics: beta_rolling=1 keeps the per-month weights in the `weights` output; pass it");
    HFREP_REQUIRE(!quality || Xtr, "predict_metrics: quality needs Xtr");
    HFREP_REQUIRE(!(post) || sigma, "predict_metrics: post needs sigma");

    EvalParams p;
    p.Xtr = Xtr; p.Xte = Xte; p.Yte = Yte; p.rf = rf; p.sigma = sigma; p.task_L = task_L; p.task_row0 = task_row0; p.W = W;
    p.ante = ante; p.post = post; p.turnover = turnover; p.metrics = metrics; p.quality = quality; p.mask = mask;
    p.weights = weights; p.beta = beta; p.norm = norm; p.status = status;
    p.n_tasks = n_tasks; p.w_stride = w_stride; p.E = c->E; p.F = c->F; p.T_train = c->T_train; p.T_test = c->T_test;
    p.window = c->window; p.want_oos = c->want_oos; p.n_oos = c->T_test - c->window - 1; p.n_pre = c->T_test - 2;
    p.acols = HFREP_MAX_L + c->F + 1;       // odd-ish stride keeps the column walks off a single bank
    p.alpha = c->alpha; p.cost_param = c->cost_param; p.phi = c->phi;

    p.rolling = c->beta_rolling ? 1 : 0;
    p.nq = p.rolling ? 4 : 1;
    // ... + the scratch of the rank-deficient path (T^T, its Householder scalars, the column order)
    p.wsz = (int)((size_t)c->window * p.acols + HFREP_MAX_L * c->F + (size_t)c->F * c->E + c->F + c->window +
                  HFREP_MAX_L * HFREP_MAX_L + 3 * HFREP_MAX_L);
    p.wsz = (p.wsz + 1) & ~1;                // keeps what follows 16-byte aligned (float4 prediction buffers)
    p.series_dbl = (int)(3 * (size_t)p.n_oos * c->F);
    if (p.series_dbl < (NTH_EVAL / 32) * 32 * WS / 2) p.series_dbl = (NTH_EVAL / 32) * 32 * WS / 2;
    p.series_dbl = (p.series_dbl + 1) & ~1;
    size_t dbl = (size_t)p.nq * p.wsz + (size_t)p.series_dbl + (NTH_EVAL / 32) * 3 * HFREP_MAX_E;
    dbl = (dbl + 1) & ~(size_t)1;           // float region starts 16-byte aligned
    p.off_float = (int)(dbl * 8);
    const size_t flt = (size_t)HFREP_MAX_E * WS + (size_t)HFREP_MAX_L * WS + (size_t)c->T_test * WS;
    const size_t smem = dbl * 8 + flt * 4 + (size_t)p.n_oos * c->E + 16;
    HFREP_REQUIRE(smem <= 227 * 1024, "predict_metrics: %zu bytes of shared memory needed (T_test=%d too long)", smem,
                  c->T_test);
    HFREP_CUDA_OK(allow_max_dynamic_smem(eval_kernel));
    // few tasks: spread each task's 167 expanding-prefix evaluations over several CTAs so the whole GPU works
    p.qsplit = 1;
    if (quality && c->want_oos) {
        const int sms = sm_count();
        while (p.qsplit < 16 && n_tasks * p.qsplit * 2 <= 2 * sms) p.qsplit *= 2;
    }
    eval_kernel<<<n_tasks * p.qsplit, NTH_EVAL, smem, (cudaStream_t)stream>>>(p);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_series_metrics(const double* series, const double* fund, const double* rf, int n, int F,
                                    double* out, void* stream) {
    HFREP_REQUIRE(series && fund && rf && out, "series_metrics: null argument");
    HFREP_REQUIRE(n >= 4 && F >= 1, "series_metrics: n=%d F=%d (the bias-corrected kurtosis needs n >= 4)", n, F);
    const int wpb = 4;
    series_metrics_kernel<<<(F + wpb - 1) / wpb, wpb * 32, 0, (cudaStream_t)stream>>>(series, fund, rf, n, F, out);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}

extern "C" int hfrep_forward(const float* W, int E, int L, const float* x, int n, int encode_only, float alpha,
                             float* out, void* stream) {
    HFREP_REQUIRE(W && x && out, "forward: null argument");
    HFREP_REQUIRE(E >= 1 && E <= HFREP_MAX_E && L >= 1 && L <= HFREP_MAX_L, "forward: E=%d L=%d", E, L);
    if (n <= 0) return HFREP_OK;
    forward_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(W, E, L, x, n, encode_only, alpha, out);
    HFREP_CUDA_OK(cudaGetLastError());
    return HFREP_OK;
}
