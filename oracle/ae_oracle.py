"""Portable CPU restatement (numpy only) of the reference's replication path.

TEST INFRASTRUCTURE (see oracle/__init__.py) — the checker for the CUDA path and the
`cpu_baseline` / `--impl reference` arm of bench.py; never imported by the product package.

Every function cites the reference lines it follows (paths relative to /root/reference).
It is validated against the reference's own files run over shims (oracle/make_golden.py →
tests/golden/*.npz; tests/test_oracle.py).  The Keras/statsmodels layer below it
(oracle/fitcore.py, pinv-OLS) is "parity unpinned" — see oracle/__init__.py.
"""
from __future__ import annotations

import numpy as np

from . import fitcore

F32, F64 = np.float32, np.float64


# ----------------------------------------------------------------------------- sklearn pieces
def minmax_fit_transform(x):
    """MinMaxScaler().fit_transform (Autoencoder_encapsulate.py:62-65, :118-119, :127-128):
    float64, per column (x - min) * scale with scale = 1/(max-min), zero range -> scale 1."""
    x = np.asarray(x, F64)
    lo, hi = x.min(axis=0), x.max(axis=0)
    rng = hi - lo
    rng = np.where(rng == 0., 1., rng)
    scale = 1. / rng
    return x * scale + (0. - lo * scale)


def r2_uniform(y, yhat):
    """sklearn r2_score, multioutput='uniform_average' (Autoencoder_encapsulate.py:109,121)."""
    y, yhat = np.asarray(y, F64), np.asarray(yhat, F64)
    num = ((y - yhat) ** 2).sum(axis=0)
    den = ((y - y.mean(axis=0)) ** 2).sum(axis=0)
    out = np.ones_like(num)
    ok = den != 0
    out[ok] = 1. - num[ok] / den[ok]
    out[(~ok) & (num != 0)] = 0.          # sklearn force_finite rule
    return float(out.mean())


def rmse_uniform(y, yhat):
    """mean_squared_error(..., squared=False) of sklearn < 1.6: per-column RMSE, then the
    uniform average (Autoencoder_encapsulate.py:113,130)."""
    y, yhat = np.asarray(y, F64), np.asarray(yhat, F64)
    return float(np.sqrt(((y - yhat) ** 2).mean(axis=0)).mean())


# ----------------------------------------------------------------------------- helper.py
def normalization(Y, X, beta, window):
    """helper.py:10-17."""
    Y, X, beta = np.asarray(Y), np.asarray(X), np.asarray(beta)
    R_hat = X @ beta
    den = np.sum((R_hat - np.mean(R_hat, axis=0)) ** 2 / (window - 1), axis=0)
    num = np.sum((Y - np.mean(Y, axis=0)) ** 2 / (window - 1), axis=0)
    return np.sqrt(num) / np.sqrt(den)


def transaction_cost(old_x, new_x, cov, param=.05):
    """helper.py:65-80."""
    s = np.sqrt(np.diag(np.asarray(cov))) * param
    d = np.asarray(old_x) - np.asarray(new_x)
    return .5 * d ** 2 * s


def price_impact(old_x, new_x, cov, param=.05, phi=.5):
    """helper.py:83-92."""
    s = np.sqrt(np.diag(np.asarray(cov))) * param
    old_x, new_x = np.asarray(old_x), np.asarray(new_x)
    d = old_x - new_x
    return phi * new_x * s * d - old_x * s * d - .5 * d ** 2 * s


def ex_post_return(ex_ante, window, strat_weight, factor_etf):
    """helper.py:112-131.  ex_ante (T,F); strat_weight (F,T,E); factor_etf (T+window,E).
    Returns (T,F): post[0]=ante[0]; post[i]=ante[i]+penalty[i-1], penalty[j] from weights j+1 vs j
    and the ddof=1 covariance of rows [j+1, j+1+window)."""
    ex_ante, factor_etf = np.asarray(ex_ante, F64), np.asarray(factor_etf, F64)
    T, Fn = ex_ante.shape
    out = np.empty_like(ex_ante)
    for f in range(Fn):
        pen = []
        for i in range(1, len(factor_etf) - window):
            cov = np.cov(factor_etf[i:i + window], rowvar=False, ddof=1)
            new_x, old_x = strat_weight[f][i], strat_weight[f][i - 1]
            pen.append((transaction_cost(old_x, new_x, cov) + price_impact(old_x, new_x, cov)).sum())
        out[0, f] = ex_ante[0, f]
        for i in range(1, T):
            out[i, f] = ex_ante[i, f] + pen[i - 1]
    return out


def random_window_starts(n_rows, n_sample, window, seed):
    """helper.py:44-62 draws `random.randint(0, T-window)` (inclusive) per sample with python's
    global Mersenne Twister; restated with an explicit seed."""
    import random
    r = random.Random(seed)
    return np.array([r.randint(0, n_rows - window) for _ in range(n_sample)], np.int32)


def random_sampling(dataset, starts, window):
    """helper.py:55-62 given the start table."""
    return np.array([dataset[s:s + window] for s in starts])


def factor_hf_split(arr, split_pos, reshape=True):
    """helper.py:133-153."""
    assert isinstance(arr, np.ndarray) and isinstance(split_pos, int) and arr.ndim == 3
    assert 0 < split_pos < arr.shape[2]
    factor, hf = arr[:, :, :split_pos].copy(), arr[:, :, split_pos:].copy()
    if reshape:
        factor = factor.reshape(-1, factor.shape[2])
        hf = hf.reshape(-1, hf.shape[2])
    return factor, hf


# ----------------------------------------------------------------------------- notebook metrics
def annualized_sharpe_ratio(ret, rf=0):
    """autoencoder_v4.ipynb:715-721."""
    ret, rf = np.asarray(ret, F64), np.asarray(rf, F64)
    return (np.mean(ret) - np.mean(rf)) / np.std(ret) * np.sqrt(12)


def tracking_error(replica, fund):
    """NOT in the reference (north_star names it): std(replica-fund, ddof=1)*sqrt(12)."""
    d = np.asarray(replica, F64) - np.asarray(fund, F64)
    return np.std(d, ddof=1) * np.sqrt(12)


def correlation(replica, fund):
    """NOT in the reference: Pearson correlation."""
    a, b = np.asarray(replica, F64), np.asarray(fund, F64)
    a, b = a - a.mean(), b - b.mean()
    return float((a * b).sum() / np.sqrt((a * a).sum() * (b * b).sum()))


def omega_ratio(r, threshold=0.):
    """autoencoder_v4.ipynb:695-705 (the (1+thr)**sqrt(1/252)-1 de-annualisation is the notebook's)."""
    thr = (threshold + 1) ** np.sqrt(1 / 252) - 1
    ex = np.asarray(r, F64) - thr
    return np.sum(ex[ex > 0]) / (-np.sum(ex[ex < 0]))


def historical_cvar(r, alpha=5):
    """autoencoder_v4.ipynb:741-757: mean of returns <= np.percentile(r, alpha) (linear interpolation)."""
    r = np.asarray(r, F64)
    return r[r <= np.percentile(r, alpha)].mean()


def ceq(ret, rf, gamma=2):
    """autoencoder_v4.ipynb:759-770."""
    ret, rf = np.asarray(ret, F64).ravel(), np.asarray(rf, F64).ravel()
    mid = np.power((1 + ret) / (1 + rf), 1 - gamma)
    return np.log(np.mean(mid)) / ((1 - gamma) / 12)


def skewness(r):
    """pandas DataFrame.skew() (bias-corrected G1), autoencoder_v4.ipynb:836."""
    r = np.asarray(r, F64)
    n = len(r)
    d = r - r.mean()
    m2, m3 = (d ** 2).mean(), (d ** 3).mean()
    return np.sqrt(n * (n - 1)) / (n - 2) * m3 / m2 ** 1.5


def kurtosis(r):
    """pandas DataFrame.kurt() (bias-corrected excess G2), autoencoder_v4.ipynb:836."""
    r = np.asarray(r, F64)
    n = len(r)
    d = r - r.mean()
    m2, m4 = (d ** 2).mean(), (d ** 4).mean()
    g2 = m4 / m2 ** 2 - 3
    return (n - 1) / ((n - 2) * (n - 3)) * ((n + 1) * g2 + 6)


# ----------------------------------------------------------------------------- class AE
class AEOracle:
    """`class AE` of Autoencoder_encapsulate.py:38-224 on plain arrays.

    x_train/x_test: (T, E) raw ETF returns; y_train/y_test: (T, F) fund returns.
    `We0`, `Wd0` (initial kernels) and `perm` (per-epoch row orders or None) are explicit
    inputs so that the CUDA path can be given the very same ones.
    """

    def __init__(self, x_train, y_train, x_test, y_test, latent_dim):
        assert len(x_train) == len(y_train) and len(y_test) == len(x_test)         # :59
        self.x_train_scaled = minmax_fit_transform(x_train)                        # :62-65
        self.x_test = np.asarray(x_test, F64)                                      # :66-67 (unscaled)
        self.y_train, self.y_test = np.asarray(y_train, F64), np.asarray(y_test, F64)
        self.L = latent_dim
        self.We = self.Wd = None

    # :72-96
    def train(self, We0, Wd0, patience=5, perm=None, cfg=None, record_grads_at=(), order="canonical"):
        cfg = cfg or fitcore.FitConfig(patience=patience)
        out = fitcore.fit(self.x_train_scaled, We0, Wd0, cfg, perm=perm, record_grads_at=record_grads_at, order=order)
        self.We, self.Wd, self.fit_out = out["We"], out["Wd"], out
        return out

    def _predict(self, x):
        return fitcore.predict(x, self.We, self.Wd)

    def model_IS_r2(self):          # :107-109
        return r2_uniform(self.x_train_scaled, self._predict(self.x_train_scaled))

    def model_IS_RMSE(self):        # :111-113
        return rmse_uniform(self.x_train_scaled, self._predict(self.x_train_scaled))

    def model_OOS_r2(self):         # :115-122
        seq = []
        for i in range(2, len(self.x_test)):
            x_real = minmax_fit_transform(self.x_test[:i])
            seq.append(r2_uniform(x_real, self._predict(x_real)))
        return seq

    def model_OOS_RMSE(self):       # :124-131
        seq = []
        for i in range(2, len(self.x_test)):
            x_real = minmax_fit_transform(self.x_test[:i])
            seq.append(rmse_uniform(x_real, self._predict(x_real)))
        return seq

    # :133-201
    def ante(self, rf, window=24, beta_index="first"):
        """rf: (T_all,) or (T_all,1) risk-free series whose LAST rows align with the test block.
        beta_index='first' reproduces Autoencoder_encapsulate.py:167 (`ae_ols_beta[0]`,
        `normalization_factor[0]`); 'rolling' uses window i (the evident intent, App. B.1)."""
        rf = np.asarray(rf, F64).reshape(-1)
        M = fitcore.encode(self.x_test, self.We)                                   # :140 float32
        n_w = len(self.x_test) - window                                             # :148
        betas, norms = [], []
        for i in range(n_w):
            X, Y = M[i:i + window], self.y_test[i:i + window]                       # :149-150
            beta = np.linalg.pinv(X.astype(F64)) @ Y                                # :151 (pinv-OLS, float64)
            betas.append(beta)
            norms.append(normalization(Y, X, beta, window))                         # :154
        Wd = self.Wd                                                                # :159 float32 (L,E)
        weights, deltas, masks = [], [], []
        for i in range(n_w):                                                        # :162
            mask = np.ones(Wd.shape[1])
            mask[(M[window + i] @ Wd) < 0] = .2                                     # :163-166
            j = 0 if beta_index == "first" else i
            w = ((betas[j].T @ Wd) * mask).T * norms[j]                             # :167  (E,F)
            deltas.append(1 - np.sum(w, axis=0))                                    # :168
            weights.append(w)
            masks.append(mask)
        weights.pop(); deltas.pop(); masks.pop()                                    # :179-180
        n = len(weights)
        self.OOS_etf = self.x_test[-n:]                                             # :183
        self.OOS_hfd = self.y_test[-n:]                                             # :184
        self.OOS_rf = rf[-n:]                                                       # :185
        ante = np.empty((n, self.y_test.shape[1]))
        for k, w in enumerate(weights):                                             # :188-190
            ante[k] = deltas[k] * self.OOS_rf[k] + np.sum(self.OOS_etf[k] * w.T, axis=1)
        self._ante = ante
        self.weights = np.array(weights)                                            # (n, E, F)
        self.masks = np.array(masks)
        self.betas, self.norms = np.array(betas), np.array(norms)
        self.latent = M
        self.window = window
        return ante

    # :203-208 + helper.py:112-131
    def post(self, factor_etf_data):
        n = len(self.weights)
        etf = np.asarray(factor_etf_data, F64)[-n - self.window:]                  # :206
        strat = np.transpose(self.weights, (2, 0, 1))                               # reshape_cab: (F, n, E)
        self._post = ex_post_return(self._ante, self.window, strat, etf)
        return self._post

    # :210-224
    def turnover(self):
        t = np.zeros(self.weights.shape[2])
        for i in range(len(self.weights) - 1):
            t += np.sum(np.abs(self.weights[i] - self.weights[i + 1]), axis=0)
        return t / (len(self.weights) / 12)

    def metrics(self, series):
        """Sharpe (notebook), tracking error and correlation (ours) of an (n,F) replica series."""
        Fn = series.shape[1]
        sh = np.array([annualized_sharpe_ratio(series[:, f], self.OOS_rf) for f in range(Fn)])
        te = np.array([tracking_error(series[:, f], self.OOS_hfd[:, f]) for f in range(Fn)])
        co = np.array([correlation(series[:, f], self.OOS_hfd[:, f]) for f in range(Fn)])
        return sh, te, co


def run_task(x_train, y_train, x_test, y_test, rf, etf_all, L, We0, Wd0, perm=None, cfg=None,
             with_oos=True, order="canonical"):
    """One complete task-fit the way the notebook sweeps it (autoencoder_v4.ipynb:158-165,
    :896-901): fit, IS/OOS quality, ante, turnover, post, metrics.  Returns a dict of arrays."""
    ae = AEOracle(x_train, y_train, x_test, y_test, L)
    fo = ae.train(We0, Wd0, perm=perm, cfg=cfg, order=order)
    out = dict(We=ae.We, Wd=ae.Wd, loss=fo["loss"], val_loss=fo["val_loss"], epochs_run=fo["epochs_run"],
               is_r2=ae.model_IS_r2(), is_rmse=ae.model_IS_RMSE())
    if with_oos:
        out["oos_r2"] = np.array(ae.model_OOS_r2())
        out["oos_rmse"] = np.array(ae.model_OOS_RMSE())
    out["ante"] = ae.ante(rf)
    out["turnover"] = ae.turnover()
    out["post"] = ae.post(etf_all)
    for name in ("ante", "post"):
        sh, te, co = ae.metrics(out[name])
        out[f"sharpe_{name}"], out[f"te_{name}"], out[f"corr_{name}"] = sh, te, co
    return out


def run_mode_w_window(etf_train, etf, hfd, rf, t, train_window, L, We0, Wd0, ols_window=24, cfg=None):
    """One task of the rolling refit (SURVEY §8d Mode W; OUR extension — the reference fits once): the reference's `AE`
    restatement called on train block rows [t-train_window, t) and test block rows [t-ols_window, t+2), `ante(window)`
    -> one replica row (for row t+1).  etf_train: the matrix the training windows are cut from."""
    lo = t - ols_window
    ae = AEOracle(etf_train[t - train_window:t], hfd[t - train_window:t], etf[lo:t + 2], hfd[lo:t + 2], L)
    ae.train(We0, Wd0, cfg=cfg)
    ante = ae.ante(np.asarray(rf).reshape(-1)[lo:t + 2], window=ols_window)
    assert ante.shape[0] == 1
    return ante[0], ae.weights[0]
