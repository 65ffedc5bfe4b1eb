/* Plain-C float32 restatement of the reference's fit — TEST INFRASTRUCTURE (see oracle/__init__.py), next to
 * oracle/fitcore.py, and the "best-effort CPU" baseline of BASELINE.md (B2).
 *
 * What it follows (paths under /root/reference; Keras 2.7 semantics restated in SURVEY.md App. A — PARITY UNPINNED for
 * that layer, Keras cannot be installed here):
 *   Autoencoder            Autoencoder_encapsulate.py:19-35   x -> lrelu(x We) -> lrelu(h Wd), no biases, alpha .2
 *   AE.train / Model.fit   Autoencoder_encapsulate.py:72-96   MSE, Nadam(), batch 48, validation tail, EarlyStopping
 *
 * Two summation orders; the file is compiled once for each (-DCANON=1 / -DCANON=0, see the Makefile), giving
 * fit_one_canonical / fit_many_canonical and fit_one_fast / fit_many_fast:
 *   1  the CANONICAL ORDER of DESIGN.md §5, the one every mapping of the CUDA fit kernel follows, so that the weight
 *      trajectory of a run is bit-identical on both sides (a LeakyReLU pre-activation within rounding of zero otherwise
 *      lets two correct float32 implementations take different branches — see DESIGN.md):
 *        zh[r][l]   Z_even + Z_odd, fused multiply-add chains over the even / the odd ETFs in ascending order, from 0
 *        zy[r][e]   the same over the even / the odd latent units
 *        dzy        lrelu'(zy) * (coef * (y - x)),  coef = 2 / (rows * E) rounded once
 *        dh[r][l]   S_even + S_odd, each a fused chain over the even / odd ETFs in ascending order
 *        dWd, dWe   (p0 + p2 + p4) + (p1 + p3 + p5), p_c = fused chain over the batch positions q = c, c+6, c+12, ...
 *        Nadam      Keras' operation order, every operation rounded on its own (no contraction)
 *      Built with -ffp-contract=off; fmaf() is the only fused operation.
 *   0  the natural loop order, built with -O3 -ffp-contract=fast (gcc may fuse and vectorise): the timed CPU baseline.
 * Built by oracle/c/Makefile into oracle/_build/; never linked by the product.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int E, L, T, n_fit, batch, epochs, patience, early_stop;
    float alpha, lr, b1, b2, eps, decay, base;
} fit_cfg;

#ifndef CANON
#define CANON 1
#endif
#if CANON
#define FIT_ONE fit_one_canonical
#define FIT_MANY fit_many_canonical
#else
#define FIT_ONE fit_one_fast
#define FIT_MANY fit_many_fast
#endif

static float lrelu(float z, float a) { return z > 0.f ? z : a * z; }

/* sum over the batch positions in canonical order: six interleaved fused chains, combined even classes first */
static float class_sum(const float* a, int sa, const float* b, int sb, int nb) {
    float p[6];
    for (int c = 0; c < 6; ++c) {
        float s = 0.f;
        for (int q = c; q < nb; q += 6) s = fmaf(a[(size_t)q * sa], b[(size_t)q * sb], s);
        p[c] = s;
    }
    return ((p[0] + p[2]) + p[4]) + ((p[1] + p[3]) + p[5]);
}

/* One fit.  x [T][E] scaled rows; We [E][L], Wd [L][E] in/out; perm [epochs][n_fit] or NULL; loss/val [epochs];
 * grad_out (optional) [2*E*L] receives dWe then dWd of global step grad_step.  Returns the number of epochs run. */
int FIT_ONE(const fit_cfg* c, const float* x, float* We, float* Wd, const int* perm, float* loss_hist, float* val_hist,
            float* grad_out, int grad_step) {
    const int E = c->E, L = c->L, B = c->batch, n_fit = c->n_fit, n_val = c->T - c->n_fit;
    const int nw = E * L, canon = CANON;
    float* buf = (float*)calloc((size_t)B * (3 * L + 3 * E) + 6 * (size_t)nw, sizeof(float));
    float *H = buf, *dZh = H + B * L, *Zh = dZh + B * L, *Zy = Zh + B * L, *dZy = Zy + B * E, *Xb = dZy + B * E;
    float *mWe = Xb + B * E, *vWe = mWe + nw, *mWd = vWe + nw, *vWd = mWd + nw, *gWe = vWd + nw, *gWd = gWe + nw;
    int* rows = (int*)malloc(sizeof(int) * B);
    float m_cache = 1.f, best = INFINITY;
    int step = 0, wait = 0, epochs_run = 0;
    for (int epoch = 0; epoch < c->epochs; ++epoch) {
        float acc = 0.f;
        int cnt = 0;
        for (int seg = 0; seg < 2; ++seg) {                         /* 0: train batches, 1: validation chunks */
            const int n = seg ? n_val : n_fit;
            float vacc = 0.f;
            int vcnt = 0;
            for (int lo = 0; lo < n; lo += B) {
                const int nb = n - lo < B ? n - lo : B;
                for (int q = 0; q < nb; ++q) rows[q] = seg ? n_fit + lo + q : (perm ? perm[(size_t)epoch * n_fit + lo + q] : lo + q);
                /* forward */
                float sse_rows = 0.f;
                for (int r = 0; r < nb; ++r) {
                    const float* xr = x + (size_t)rows[r] * E;
                    memcpy(Xb + (size_t)r * E, xr, sizeof(float) * E);
                    for (int l = 0; l < L; ++l) {
                        float z = 0.f;
                        if (canon) {
                            float z0 = 0.f, z1 = 0.f;
                            for (int e = 0; e < E; e += 2) z0 = fmaf(xr[e], We[e * L + l], z0);
                            for (int e = 1; e < E; e += 2) z1 = fmaf(xr[e], We[e * L + l], z1);
                            z = z0 + z1;
                        } else for (int e = 0; e < E; ++e) z += xr[e] * We[e * L + l];
                        Zh[r * L + l] = z;
                        H[r * L + l] = lrelu(z, c->alpha);
                    }
                    float row = 0.f;
                    for (int e = 0; e < E; ++e) {
                        float z = 0.f;
                        if (canon) {
                            float z0 = 0.f, z1 = 0.f;
                            for (int l = 0; l < L; l += 2) z0 = fmaf(H[r * L + l], Wd[l * E + e], z0);
                            for (int l = 1; l < L; l += 2) z1 = fmaf(H[r * L + l], Wd[l * E + e], z1);
                            z = z0 + z1;
                        } else for (int l = 0; l < L; ++l) z += H[r * L + l] * Wd[l * E + e];
                        Zy[r * E + e] = z;
                        const float d = lrelu(z, c->alpha) - xr[e];
                        row += d * d;
                    }
                    sse_rows += row / (float)E;                    /* mean over the last axis ... */
                }
                const float batch_loss = sse_rows / (float)nb;     /* ... then SUM_OVER_BATCH_SIZE (App. A.2) */
                if (seg) { vacc += batch_loss * nb; vcnt += nb; continue; }
                acc += batch_loss * nb; cnt += nb;
                /* backward */
                const float coef = 2.f / (float)(nb * E);
                for (int r = 0; r < nb; ++r) {
                    const float* xr = Xb + (size_t)r * E;
                    for (int e = 0; e < E; ++e) {
                        const float z = Zy[r * E + e];
                        const float g = canon ? coef * (lrelu(z, c->alpha) - xr[e])
                                              : 2.f * (lrelu(z, c->alpha) - xr[e]) / (float)(nb * E);
                        dZy[r * E + e] = z > 0.f ? g : c->alpha * g;
                    }
                    for (int l = 0; l < L; ++l) {
                        float g;
                        if (canon) {
                            float s0 = 0.f, s1 = 0.f;
                            for (int e = 0; e < E; e += 2) s0 = fmaf(dZy[r * E + e], Wd[l * E + e], s0);
                            for (int e = 1; e < E; e += 2) s1 = fmaf(dZy[r * E + e], Wd[l * E + e], s1);
                            g = s0 + s1;
                        } else {
                            g = 0.f;
                            for (int e = 0; e < E; ++e) g += dZy[r * E + e] * Wd[l * E + e];
                        }
                        dZh[r * L + l] = Zh[r * L + l] > 0.f ? g : c->alpha * g;
                    }
                }
                if (canon) {
                    for (int l = 0; l < L; ++l)
                        for (int e = 0; e < E; ++e) gWd[l * E + e] = class_sum(H + l, L, dZy + e, E, nb);
                    for (int e = 0; e < E; ++e)
                        for (int l = 0; l < L; ++l) gWe[e * L + l] = class_sum(Xb + e, E, dZh + l, L, nb);
                } else {
                    memset(gWe, 0, sizeof(float) * 2 * nw);        /* gWe and gWd are adjacent */
                    for (int r = 0; r < nb; ++r) {
                        const float* xr = Xb + (size_t)r * E;
                        for (int l = 0; l < L; ++l)
                            for (int e = 0; e < E; ++e) gWd[l * E + e] += H[r * L + l] * dZy[r * E + e];
                        for (int e = 0; e < E; ++e)
                            for (int l = 0; l < L; ++l) gWe[e * L + l] += xr[e] * dZh[r * L + l];
                    }
                }
                if (grad_out && step == grad_step) memcpy(grad_out, gWe, sizeof(float) * 2 * nw);
                /* Nadam (App. A.4), float32 scalars, powers as correctly rounded float32 */
                const float t = (float)(step + 1), t1 = (float)(step + 2);
                const float mu_t = c->b1 * (1.f - .5f * (float)pow((double)c->base, (double)(c->decay * t)));
                const float mu_t1 = c->b1 * (1.f - .5f * (float)pow((double)c->base, (double)(c->decay * t1)));
                const float ms_new = m_cache * mu_t, ms_next = ms_new * mu_t1;
                const float vden = 1.f - (float)pow((double)c->b2, (double)t);
                const float om_new = 1.f - ms_new, om_next = 1.f - ms_next, om_mu = 1.f - mu_t;
                const float omb1 = 1.f - c->b1, omb2 = 1.f - c->b2;
                for (int which = 0; which < 2; ++which) {
                    float *w = which ? Wd : We, *g = which ? gWd : gWe, *m = which ? mWd : mWe, *v = which ? vWd : vWe;
                    for (int i = 0; i < nw; ++i) {
                        const float gp = g[i] / om_new;
                        const float t1m = c->b1 * m[i], t2m = omb1 * g[i];
                        m[i] = t1m + t2m;
                        const float mp = m[i] / om_next;
                        const float gg = g[i] * g[i];
                        const float t1v = c->b2 * v[i], t2v = omb2 * gg;
                        v[i] = t1v + t2v;
                        const float vp = v[i] / vden;
                        const float a1 = om_mu * gp, a2 = mu_t1 * mp;
                        const float mbar = a1 + a2;
                        const float num = c->lr * mbar;
                        w[i] = w[i] - num / (sqrtf(vp) + c->eps);
                    }
                }
                m_cache = ms_new;
                ++step;
            }
            if (seg) val_hist[epoch] = vcnt ? vacc / (float)vcnt : NAN;
        }
        loss_hist[epoch] = acc / (float)cnt;
        epochs_run = epoch + 1;
        if (c->early_stop && n_val > 0) {                          /* EarlyStopping.on_epoch_end (App. A.6); without */
            ++wait;                                                /* validation rows Keras warns and never stops    */
            if (val_hist[epoch] < best) { best = val_hist[epoch]; wait = 0; }
            if (wait >= c->patience && epoch > 0) break;
        }
    }
    free(rows);
    free(buf);
    return epochs_run;
}

/* n independent fits on all cores (OpenMP over tasks): task i uses x (shared), We/Wd at stride 2*E*Lmax, its own L. */
void FIT_MANY(const fit_cfg* c, int n, const int* task_L, const float* x, float* W, int w_stride, float* loss_hist,
              float* val_hist, int* epochs_run) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < n; ++i) {
        fit_cfg ci = *c;
        ci.L = task_L[i];
        float* w = W + (size_t)i * w_stride;
        epochs_run[i] = FIT_ONE(&ci, x, w, w + ci.E * ci.L, NULL, loss_hist + (size_t)i * c->epochs,
                                val_hist + (size_t)i * c->epochs, NULL, 0);
    }
}
