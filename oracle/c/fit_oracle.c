/* Plain-C float32 restatement of the reference's fit — TEST INFRASTRUCTURE (see oracle/__init__.py), the second,
 * independent restatement next to oracle/fitcore.py and the "best-effort CPU" baseline of BASELINE.md (B2).
 *
 * What it follows (paths under /root/reference; Keras 2.7 semantics restated in SURVEY.md App. A — PARITY UNPINNED for
 * that layer, Keras cannot be installed here):
 *   Autoencoder            Autoencoder_encapsulate.py:19-35   x -> lrelu(x We) -> lrelu(h Wd), no biases, alpha .2
 *   AE.train / Model.fit   Autoencoder_encapsulate.py:72-96   MSE, Nadam(), batch 48, validation tail, EarlyStopping
 * Built by oracle/c/Makefile (gcc -O2 -fopenmp) into oracle/_build/libfit_oracle.so; never linked by the product.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    int E, L, T, n_fit, batch, epochs, patience, early_stop;
    float alpha, lr, b1, b2, eps, decay, base;
} fit_cfg;

static float lrelu(float z, float a) { return z > 0.f ? z : a * z; }

/* One fit.  x [T][E] scaled rows; We [E][L], Wd [L][E] in/out; perm [epochs][n_fit] or NULL; loss/val [epochs].
 * Returns the number of epochs run. */
int fit_one(const fit_cfg* c, const float* x, float* We, float* Wd, const int* perm, float* loss_hist, float* val_hist) {
    const int E = c->E, L = c->L, B = c->batch, n_fit = c->n_fit, n_val = c->T - c->n_fit;
    const int nw = E * L;
    float* buf = (float*)calloc((size_t)B * (2 * L + 2 * E) + 6 * (size_t)nw, sizeof(float));
    float *H = buf, *dZh = H + B * L, *Zy = dZh + B * L, *dZy = Zy + B * E;
    float *mWe = dZy + B * E, *vWe = mWe + nw, *mWd = vWe + nw, *vWd = mWd + nw, *gWe = vWd + nw, *gWd = gWe + nw;
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
                    for (int l = 0; l < L; ++l) {
                        float z = 0.f;
                        for (int e = 0; e < E; ++e) z += xr[e] * We[e * L + l];
                        H[r * L + l] = lrelu(z, c->alpha);
                    }
                    float row = 0.f;
                    for (int e = 0; e < E; ++e) {
                        float z = 0.f;
                        for (int l = 0; l < L; ++l) z += H[r * L + l] * Wd[l * E + e];
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
                for (int r = 0; r < nb; ++r) {
                    const float* xr = x + (size_t)rows[r] * E;
                    for (int e = 0; e < E; ++e) {
                        const float z = Zy[r * E + e];
                        const float g = 2.f * (lrelu(z, c->alpha) - xr[e]) / (float)(nb * E);
                        dZy[r * E + e] = z > 0.f ? g : c->alpha * g;
                    }
                    for (int l = 0; l < L; ++l) {
                        float g = 0.f;
                        for (int e = 0; e < E; ++e) g += dZy[r * E + e] * Wd[l * E + e];
                        dZh[r * L + l] = H[r * L + l] > 0.f ? g : c->alpha * g;
                    }
                }
                memset(gWe, 0, sizeof(float) * 2 * nw);            /* gWe and gWd are adjacent */
                for (int r = 0; r < nb; ++r) {
                    const float* xr = x + (size_t)rows[r] * E;
                    for (int l = 0; l < L; ++l)
                        for (int e = 0; e < E; ++e) gWd[l * E + e] += H[r * L + l] * dZy[r * E + e];
                    for (int e = 0; e < E; ++e)
                        for (int l = 0; l < L; ++l) gWe[e * L + l] += xr[e] * dZh[r * L + l];
                }
                /* Nadam (App. A.4), float32 scalars, powers as correctly rounded float32 */
                const float t = (float)(step + 1), t1 = (float)(step + 2);
                const float mu_t = c->b1 * (1.f - .5f * (float)pow((double)c->base, (double)(c->decay * t)));
                const float mu_t1 = c->b1 * (1.f - .5f * (float)pow((double)c->base, (double)(c->decay * t1)));
                const float ms_new = m_cache * mu_t, ms_next = ms_new * mu_t1;
                const float vden = 1.f - (float)pow((double)c->b2, (double)t);
                for (int which = 0; which < 2; ++which) {
                    float *w = which ? Wd : We, *g = which ? gWd : gWe, *m = which ? mWd : mWe, *v = which ? vWd : vWe;
                    for (int i = 0; i < nw; ++i) {
                        const float gp = g[i] / (1.f - ms_new);
                        m[i] = c->b1 * m[i] + (1.f - c->b1) * g[i];
                        const float mp = m[i] / (1.f - ms_next);
                        v[i] = c->b2 * v[i] + (1.f - c->b2) * g[i] * g[i];
                        const float vp = v[i] / vden;
                        w[i] -= c->lr * ((1.f - mu_t) * gp + mu_t1 * mp) / (sqrtf(vp) + c->eps);
                    }
                }
                m_cache = ms_new;
                ++step;
            }
            if (seg) val_hist[epoch] = vcnt ? vacc / (float)vcnt : NAN;
        }
        loss_hist[epoch] = acc / (float)cnt;
        epochs_run = epoch + 1;
        if (c->early_stop) {                                       /* EarlyStopping.on_epoch_end (App. A.6) */
            ++wait;
            if (val_hist[epoch] < best) { best = val_hist[epoch]; wait = 0; }
            if (wait >= c->patience && epoch > 0) break;
        }
    }
    free(rows);
    free(buf);
    return epochs_run;
}

/* n independent fits on all cores (OpenMP over tasks): task i uses x (shared), We/Wd at stride 2*E*Lmax, its own L. */
void fit_many(const fit_cfg* c, int n, const int* task_L, const float* x, float* W, int w_stride, float* loss_hist,
              float* val_hist, int* epochs_run) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < n; ++i) {
        fit_cfg ci = *c;
        ci.L = task_L[i];
        float* w = W + (size_t)i * w_stride;
        epochs_run[i] = fit_one(&ci, x, w, w + ci.E * ci.L, NULL, loss_hist + (size_t)i * c->epochs, val_hist + (size_t)i * c->epochs);
    }
}
