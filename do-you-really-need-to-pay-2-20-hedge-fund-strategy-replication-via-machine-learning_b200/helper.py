"""Drop-in for the reference's ``helper.py`` — same names, argument order and defaults
(helper.py:10-162) — with the array work done by the CUDA library (kernel (a) window
gather, rolling volatility + ex-post penalty kernels, series metrics).  Importing it
requires the built library; there is no CPU fallback for the GPU-backed functions.

`normalization`, `transaction_cost` and `price_impact` are one-line formulas on a handful of
values; as stand-alone helpers they stay host arithmetic (inside ``AE.ante``/``AE.post`` the
same formulas run fused in kernel (c) / the ex-post kernel).
"""
from __future__ import annotations

import os
import pickle
import random

import numpy as np
import pandas as pd
import torch

from . import engine
from ._lib import lib, check, ptr, stream_ptr

_STATE = {"seed": 123, "draws": 0}


def normalization(Y, X, beta, window):
    """helper.py:10-17: sqrt(var(Y)) / sqrt(var(X beta)), ddof = window-1 ... per fund."""
    Y, X, beta = np.array(Y), np.array(X), np.array(beta)
    R_hat = X @ beta
    denominator = np.sum((R_hat - np.mean(R_hat, axis=0)) ** 2 / (window - 1), axis=0)
    numerator = np.sum((Y - np.mean(Y, axis=0)) ** 2 / (window - 1), axis=0)
    return np.sqrt(numerator) / np.sqrt(denominator)


def read_csv(loc, date=True):
    """helper.py:18-23."""
    df = pd.read_csv(loc)
    if date:
        df['Date'] = pd.to_datetime(df['Date'])
        df.set_index('Date', inplace=True)
    return df


def dic_read(loc):
    """helper.py:26-29."""
    with open(loc, "rb") as f:
        return pickle.load(f)


def set_seed(seed_value=123):
    """helper.py:32-41 seeds python/numpy/TensorFlow and pins TF to one thread.  Here it seeds
    python's and numpy's global generators the same way and restarts the Philox streams from which
    ``AE.train`` draws initial kernels and epoch permutations."""
    os.environ['PYTHONHASHSEED'] = str(seed_value)
    np.random.seed(seed_value)
    random.seed(seed_value)
    _STATE["seed"], _STATE["draws"] = int(seed_value), 0


def next_stream_seed() -> int:
    """Seed for the next model built after set_seed (successive models get distinct streams,
    as successive Keras models do)."""
    s = _STATE["seed"] + 1000003 * _STATE["draws"]
    _STATE["draws"] += 1
    return s


def _as_words(a: np.ndarray):
    """View a float32/float64 C-contiguous 2-D array as 32-bit words so kernel (a) moves it bit-exactly."""
    a = np.ascontiguousarray(a)
    if a.dtype == np.float32:
        return a, 1
    if a.dtype != np.float64:
        a = a.astype(np.float64)
    return a.view(np.float32), 2


def random_sampling(dataset, n_sample, window):
    """helper.py:44-62: n_sample windows dataset[s:s+window] with s = random.randint(0, T-window)
    (inclusive, python's global generator) -> array (n_sample, window, cols).  Starts are drawn on
    the host exactly as the reference draws them; the gather runs on the GPU (kernel (a))."""
    isinstance(dataset, np.ndarray)
    starts = np.array([random.randint(0, dataset.shape[0] - window) for _ in range(n_sample)], np.int32)
    data = np.asarray(dataset)
    if data.ndim != 2:
        raise ValueError("random_sampling expects a 2-D array")
    words, k = _as_words(data)
    dev = engine._dev()
    R = engine.to_device(words, dev)
    out = engine.window_gather(R, starts, window)
    res = out.cpu().numpy()
    if k == 2:
        res = res.view(np.float64)
    return res.astype(data.dtype, copy=False)


def transaction_cost(old_x, new_x, covMatrix, param=0.05):
    """helper.py:65-80."""
    isinstance(param, float)
    sigma = np.sqrt(np.diag(np.asarray(covMatrix))) * param
    delta_x = np.array(old_x) - np.array(new_x)
    return 0.5 * delta_x ** 2 * sigma


def price_impact(old_x, new_x, covMatrix, param=0.05, phi=0.5):
    """helper.py:83-92."""
    isinstance(param, float)
    isinstance(phi, float)
    sigma = np.sqrt(np.diag(np.asarray(covMatrix))) * param
    old_x, new_x = np.array(old_x), np.array(new_x)
    delta_x = old_x - new_x
    return phi * new_x * sigma * delta_x - old_x * sigma * delta_x - 0.5 * delta_x ** 2 * sigma


def reshape_cab(df_list):
    """helper.py:94-110: list of A frames (B x C) -> C frames (A x B); a stride change, done in one
    transpose instead of A*C row appends."""
    assert isinstance(df_list, list)
    assert isinstance(df_list[0], pd.DataFrame)
    cube = np.stack([df.values for df in df_list])            # (A, B, C)
    cols, rows = df_list[0].columns, df_list[0].index
    out = []
    for i in range(cube.shape[2]):
        out.append(pd.DataFrame(cube[:, :, i], index=[cols[i]] * cube.shape[0], columns=rows))
    return out


def ex_post_return(ex_ante, window, strat_weight, factor_etf):
    """helper.py:112-131 on the GPU: rolling ddof=1 volatilities of factor_etf (rows [i, i+window),
    i = 1..T-1) then post[0] = ante[0], post[t] = ante[t] + sum_e(tc + pi)(w[t-1], w[t])."""
    assert isinstance(ex_ante, pd.DataFrame)
    assert isinstance(factor_etf, pd.DataFrame)
    T, F = ex_ante.shape
    E = factor_etf.shape[1]
    dev = engine._dev()
    with torch.cuda.device(dev):
        etf = engine.to_device(np.asarray(factor_etf.values, np.float64), dev)
        starts, _ = engine.window_index("expost", len(factor_etf), window)
        n_pen = len(starts)                                       # len(factor_etf) - window - 1
        if n_pen < T - 1:
            raise IndexError("list index out of range")           # what helper.py:129 would raise
        sigma = engine.rolling_sigma(etf, torch.from_numpy(starts).to(dev), window)
        w = np.stack([np.asarray(sw.values if hasattr(sw, "values") else sw, np.float64)[:T] for sw in strat_weight])
        wd = engine.to_device(w, dev)                             # (F, T, E)
        ante = engine.to_device(np.asarray(ex_ante.values, np.float64), dev)
        post = torch.empty_like(ante)
        check(lib.hfrep_ex_post_return(ptr(ante), ptr(wd), ptr(sigma), T, F, E, 0.05, 0.5, ptr(post), stream_ptr()))
        res = post.cpu().numpy()
    return pd.DataFrame(res.T, columns=ex_ante.index, index=ex_ante.columns).T


def factor_hf_split(arr, split_pos, reshape=True):
    """helper.py:133-153: (n, T, C) -> factor (n, T, split_pos) and hf (n, T, C-split_pos), flattened
    to (n*T, .) when reshape.  Two column-range gathers of kernel (a)."""
    assert isinstance(arr, np.ndarray)
    assert isinstance(split_pos, int)
    assert arr.ndim == 3
    assert 0 < split_pos < arr.shape[2]
    n, T, Cc = arr.shape
    words, k = _as_words(arr.reshape(n * T, Cc))
    dev = engine._dev()
    R = engine.to_device(words, dev)
    starts = (np.arange(n, dtype=np.int64) * T).astype(np.int32)
    fa = engine.window_gather(R, starts, T, 0, split_pos * k).cpu().numpy()
    hf = engine.window_gather(R, starts, T, split_pos * k, (Cc - split_pos) * k).cpu().numpy()
    if k == 2:
        fa, hf = fa.view(np.float64), hf.view(np.float64)
    fa, hf = fa.astype(arr.dtype, copy=False), hf.astype(arr.dtype, copy=False)
    if reshape:
        fa = fa.reshape(n * T, split_pos)
        hf = hf.reshape(n * T, Cc - split_pos)
    return fa, hf


def dic_save(dic, loc):
    """helper.py:155-162."""
    with open(loc, 'wb') as f:
        pickle.dump(dic, f)
    output = dic_read(loc)
    print('stored dictionary:\n')
    print(output)


# ---- additive API (not in the reference's helper.py) ------------------------------
def series_statistics(series, fund, rf):
    """GPU statistics of an (n, F) return block: dict name -> (F,) for _lib.STAT_NAMES (Sharpe, tracking error,
    correlation, Omega(0 %), Omega(10 %), cVaR(5 %), CEQ(2/5/10), skewness, kurtosis; autoencoder_v4.ipynb:695-772, :836)."""
    from ._lib import N_STATS, STAT_NAMES
    s = np.asarray(series, np.float64)
    y = np.asarray(fund, np.float64)
    if s.ndim == 1:
        s, y = s[:, None], y[:, None]
    r = np.asarray(rf, np.float64).reshape(-1)
    dev = engine._dev()
    with torch.cuda.device(dev):
        sd, yd, rd = (engine.to_device(a, dev) for a in (s, y, r))
        out = torch.empty((N_STATS, s.shape[1]), dtype=torch.float64, device=dev)
        check(lib.hfrep_series_metrics(ptr(sd), ptr(yd), ptr(rd), s.shape[0], s.shape[1], ptr(out), stream_ptr()))
        o = out.cpu().numpy()
    return dict(zip(STAT_NAMES, o))


def replica_metrics(series, fund, rf):
    """(sharpe, tracking_error, correlation), each (F,) — see series_statistics."""
    st = series_statistics(series, fund, rf)
    return st["sharpe"], st["tracking_error"], st["correlation"]


# Fama-French factor frames, as the notebook keeps them in globals (autoencoder_v4.ipynb cells "three_factor = ...",
# "five_factor = ..."): assign them (e.g. from load_ff_factors) before data_analysis(real_data=True), or pass them in.
three_factor = None
five_factor = None


def load_ff_factors(path, start='1994-04-30', end='2022-04-30'):
    """The notebook's factor cells: daily Mkt-RF / SMB / HML in per cent -> calendar-month sums -> log(x/100 + 1) ->
    rows start..end.  Used for both files: the notebook reads the same three columns from the 5-factor file."""
    f = pd.read_csv(path, usecols=['Date', 'Mkt-RF', 'SMB', 'HML'])
    f['Date'] = pd.to_datetime(f['Date'], format='%Y%m%d')
    f = f.set_index('Date')
    try:
        f = f.resample('ME').sum()
    except ValueError:                       # pandas < 2.2 spells month-end 'M'
        f = f.resample('M').sum()
    return np.log(f / 100 + 1).loc[start:end, :]


def OLS_alpha(ret, X):
    """Intercept of ret ~ const + X (the notebook's sm.OLS(ret, sm.add_constant(X)).fit().params[0]); values only,
    like statsmodels (no index alignment)."""
    y = np.asarray(ret, np.float64).reshape(-1)
    Z = np.asarray(X, np.float64)
    Z = Z.reshape(len(y), -1)
    return float(np.linalg.lstsq(np.column_stack([np.ones(len(y)), Z]), y, rcond=None)[0][0])


def grstest(ret_mat, factor_mat):
    """Gibbons-Ross-Shanken test, restated from the R function of the notebook (cell "grstest<-function").
    Returns [[F], [p]] like the R matrix (F: N x N for N test assets; the notebook passes one)."""
    from scipy import stats as _st
    R = np.asarray(ret_mat, np.float64)
    R = R.reshape(len(R), -1)
    Fm = np.asarray(factor_mat, np.float64)
    Fm = Fm.reshape(len(R), -1)
    T, N = R.shape
    K = Fm.shape[1]
    dat = np.column_stack([np.ones(T), Fm])
    b = np.linalg.solve(dat.T @ dat, dat.T @ R)              # (K+1, N)
    e = R - dat @ b
    sigma = e.T @ e / (T - K - 1)
    alpha = b[0].reshape(N, 1)
    mean = Fm.mean(axis=0, keepdims=True)
    omega = (Fm - mean).T @ (Fm - mean) / (T - 1)
    tem1 = alpha.T @ np.linalg.solve(sigma, alpha)
    tem2 = 1 + mean @ np.linalg.solve(omega, mean.T)
    F = (T / N) * ((T - N - K) / (T - K - 1)) * (tem1 / tem2)
    return np.vstack([F, _st.f.sf(F, N, T - N - K)])


def hktest(rt, rb):
    """Huberman-Kandel spanning test, restated from the R function of the notebook (cell "hktest <- function",
    pracma::mldivide / inv / eig, corpcor::pseudoinverse).  Returns [[F], [p]]."""
    from scipy import stats as _st
    Rt = np.asarray(rt, np.float64)
    Rt = Rt.reshape(len(Rt), -1)
    Rb = np.asarray(rb, np.float64)
    Rb = Rb.reshape(len(Rt), -1)
    Tn, N = Rt.shape
    K = Rb.shape[1]
    A = np.vstack([np.concatenate([[1.], np.zeros(K)]), np.concatenate([[0.], -np.ones(K)])])
    Cm = np.vstack([np.zeros((1, N)), -np.ones((1, N))])
    X = np.column_stack([np.ones(Tn), Rb])
    B = np.linalg.lstsq(X, Rt, rcond=None)[0]
    Theta = A @ B - Cm
    e = Rt - X @ B
    Sigma = np.atleast_2d(np.cov(e, rowvar=False))
    H = Theta @ np.linalg.inv(Sigma) @ Theta.T
    mu1 = Rb.mean(axis=0, keepdims=True)
    V11i = np.linalg.pinv(np.atleast_2d(np.cov(Rb, rowvar=False)))
    a1 = (mu1 @ V11i @ mu1.T).item()
    b1 = float(np.sum(V11i @ mu1.T))
    c1 = float(np.sum(V11i))
    G = np.array([[1 + a1, b1], [b1, c1]])
    lam = np.linalg.eigvals(H @ np.linalg.inv(G))
    Ui = float(np.real(np.prod(1 + lam)))
    if N == 1:
        Ftest = (Tn - K - 1) * (Ui - 1) / 2
        pval = _st.f.sf(Ftest, 2, Tn - K - 1)
    else:
        Ftest = (Tn - K - N) * (np.sqrt(Ui) - 1) / N
        pval = _st.f.sf(Ftest, 2 * N, 2 * (Tn - N - K))
    return np.array([[Ftest], [pval]])


def data_analysis(df, name, rf=None, start=None, end=None, span=None, real_data=True, three_factor=None, five_factor=None):
    """The notebook's report table (autoencoder_v4.ipynb:774-849).  The per-series statistics run on the GPU
    (hfrep_series_metrics); real_data=True adds, as in the notebook, the Fama-French 3- and "5"-factor alphas (the
    notebook reads the same three columns from both files) and the GRS / HK tests against `span` (or the other columns),
    restated on the host from the notebook's statsmodels call and R functions.  The factor frames come from the module
    globals `three_factor` / `five_factor` (the notebook's globals) unless passed in."""
    vals = np.asarray(df, np.float64)
    rfv = np.zeros(len(vals)) if rf is None else np.asarray(rf, np.float64).reshape(-1)
    st = series_statistics(vals, vals, rfv)
    cols = {'Skewness': st["skewness"], 'Kurtosis': st["kurtosis"], 'Omega_ratio(0%)': st["omega0"],
            'Omega_ratio(10%)': st["omega10"], 'cVaR(95%)': st["cvar5"], 'CEQ Gamma=2': st["ceq2"],
            'CEQ Gamma=5': st["ceq5"], 'CEQ Gamma=10': st["ceq10"], 'Annualized_Sharpe': st["sharpe"]}
    if real_data:
        g = globals()
        f3 = three_factor if three_factor is not None else g.get("three_factor")
        f5 = five_factor if five_factor is not None else g.get("five_factor")
        if f3 is None or f5 is None:
            raise ValueError("data_analysis(real_data=True) needs the Fama-French factor frames: set helper.three_factor / "
                             "helper.five_factor (see load_ff_factors) or pass three_factor= / five_factor=")
        cut = lambda x: x.loc[start:end, ] if (start and end and hasattr(x, "loc")) else x
        f3v, f5v = np.asarray(cut(f3), np.float64), np.asarray(cut(f5), np.float64)
        spanv = None if span is None else np.asarray(cut(span), np.float64)
        a3, a5, grsf, grsp, hkf, hkp = [], [], [], [], [], []
        for j in range(vals.shape[1]):
            a3.append(OLS_alpha(vals[:, j], f3v))
            a5.append(OLS_alpha(vals[:, j], f5v))
            bench = spanv if spanv is not None else np.delete(vals, j, axis=1)
            hk, grs = hktest(vals[:, j], bench), grstest(vals[:, j], bench)
            hkf.append(hk[0][0]); hkp.append(round(float(hk[1][0]), 6))
            grsf.append(grs[0][0]); grsp.append(round(float(grs[1][0]), 6))
        cols.update({'FF3F_alpha': a3, 'FF5F_alpha': a5, 'GRS_testF': grsf, 'HK_testF': hkf, 'GRS_test_pval': grsp,
                     'HK_test_pval': hkp})
    stats = pd.DataFrame(cols)
    stats.index = name
    return stats


def annualized_sharpe_ratio(ret, rf=0):
    """autoencoder_v4.ipynb:715-721 for one series."""
    ret = np.asarray(ret, np.float64).reshape(-1)
    rfv = np.broadcast_to(np.asarray(rf, np.float64).reshape(-1), ret.shape) if np.ndim(rf) else np.full(ret.shape, float(rf))
    return float(replica_metrics(ret, ret, rfv)[0][0])


def tracking_error(replica, fund):
    r = np.asarray(replica, np.float64).reshape(-1)
    return float(replica_metrics(r, np.asarray(fund, np.float64).reshape(-1), np.zeros_like(r))[1][0])


def correlation(replica, fund):
    r = np.asarray(replica, np.float64).reshape(-1)
    return float(replica_metrics(r, np.asarray(fund, np.float64).reshape(-1), np.zeros_like(r))[2][0])
