"""float32 numpy restatement of the Keras 2.7 pieces the reference's fit relies on.

TEST INFRASTRUCTURE (see oracle/__init__.py).  **Parity unpinned**: Keras/TensorFlow are
not installable here and the reference has no tests; what follows restates Keras 2.7
(`optimizer_v2`) semantics as listed in SURVEY.md App. A.1-A.6, each item tied to the
reference call site that relies on it.

    Dense(use_bias=False) + LeakyReLU(alpha=.2)      Autoencoder_encapsulate.py:23-30
    MeanSquaredError()                               Autoencoder_encapsulate.py:81
    Nadam() defaults                                 Autoencoder_encapsulate.py:80
    fit(epochs, batch_size, validation_split)        Autoencoder_encapsulate.py:83-96
    EarlyStopping(monitor='val_loss', patience)      Autoencoder_encapsulate.py:90-94
    predict()                                        Autoencoder_encapsulate.py:108,112,120,129,140

All arithmetic is float32 (Keras casts inputs and keeps variables in float32).

Summation order.  Float32 sums depend on their order, and this fit has a discrete element: a LeakyReLU
pre-activation that passes within rounding of zero takes one branch or the other depending on the last bit of a
22-term sum, which changes that row's gradient five-fold for one step (measured on the reference's data: L = 8,
step 1993, row 87, unit 6, z = +4.1e-8 against -3.9e-8; the validation loss of that epoch then differs by 3.2e-5
and the difference decays over the next epochs).  Two correct float32 implementations therefore agree to 1e-6 on
most runs and to 3e-5 on some.  To make parity exact rather than lucky, `fit(order="canonical")` fixes the order
of every sum that feeds the weights (the CANONICAL ORDER) — the CUDA fit kernel and the C restatement
(oracle/c/fit_oracle.c) follow the same order, so the three produce bit-identical weight trajectories.
`order="blas"` leaves the sums to numpy's BLAS (what Keras/Eigen effectively does; the timed CPU baseline).
"""
from __future__ import annotations

import dataclasses
import numpy as np

F32 = np.float32


@dataclasses.dataclass
class FitConfig:
    """Literals of Autoencoder_encapsulate.py:79-96 and Nadam()'s defaults."""
    epochs: int = 1000            # :86
    batch_size: int = 48          # :88
    validation_split: float = .25  # :89
    patience: int = 5             # :72 / :92
    early_stop: bool = True
    alpha: float = .2             # :25, :29
    lr: float = 1e-3              # Nadam() defaults (Keras 2.7)
    beta_1: float = .9
    beta_2: float = .999
    epsilon: float = 1e-7
    schedule_decay: float = .004
    decay_base: float = .96


def glorot_limit(fan_in: int, fan_out: int) -> float:
    """GlorotUniform limit, Dense's default kernel initialiser (App. A.1)."""
    return float(np.sqrt(6.0 / (fan_in + fan_out)))


def lrelu(z: np.ndarray, alpha: float) -> np.ndarray:
    """LeakyReLU(alpha): z if z > 0 else alpha*z (Autoencoder_encapsulate.py:25,29)."""
    return np.where(z > 0, z, F32(alpha) * z).astype(F32, copy=False)


def forward(x: np.ndarray, We: np.ndarray, Wd: np.ndarray, alpha: float):
    """Autoencoder.call (Autoencoder_encapsulate.py:32-35): returns zh, h, zy, y."""
    zh = x @ We
    h = lrelu(zh, alpha)
    zy = h @ Wd
    y = lrelu(zy, alpha)
    return zh, h, zy, y


def encode(x: np.ndarray, We: np.ndarray, alpha: float = .2) -> np.ndarray:
    """encoder.predict (Autoencoder_encapsulate.py:140)."""
    return lrelu(np.asarray(x, F32) @ We, alpha)


def predict(x: np.ndarray, We: np.ndarray, Wd: np.ndarray, alpha: float = .2) -> np.ndarray:
    """Model.predict: float32 forward pass (App. A.5)."""
    return forward(np.asarray(x, F32), We, Wd, alpha)[3]


def fma32(a, b, c):
    """RN_float32(a*b + c) with ONE rounding — the fused multiply-add of the CUDA kernel (`fma.rn.f32`) and of C's
    fmaf(), emulated in float64: the product of two float32 is exact in float64; the float64 sum is rounded once more
    when cast to float32, which differs from the single rounding only when the float64 sum lands exactly on the midpoint
    of two float32 neighbours while the exact sum does not — detected with the exact rounding error of the addition
    (TwoSum) and repaired by nudging the sum off the midpoint in the error's direction."""
    p = np.asarray(a, np.float64) * np.asarray(b, np.float64)
    c64 = np.asarray(c, np.float64)
    s = p + c64
    bb = s - p
    err = (p - (s - bb)) + (c64 - bb)
    tie = ((s.view(np.uint64) & np.uint64(0x1FFFFFFF)) == np.uint64(0x10000000)) & (err != 0)
    if tie.any():
        s = np.where(tie, np.nextafter(s, np.where(err > 0, np.inf, -np.inf)), s)
    return s.astype(F32)


def chain(a, b, axis_terms):
    """Fused chain  acc = fma(a_k, b_k, acc), k in axis_terms order, from 0.  a: (n, K), b: (K, m) -> (n, m)."""
    acc = np.zeros((a.shape[0], b.shape[1]), F32)
    for k in axis_terms:
        acc = fma32(a[:, k:k + 1], b[k:k + 1, :], acc)
    return acc


def class_sum(a, b):
    """Canonical sum over batch positions (rows of a and b):  (p0 + p2 + p4) + (p1 + p3 + p5),  p_c = fused chain over the
    positions q = c, c+6, c+12, ...  a: (nb, n), b: (nb, m) -> a^T b (n, m)."""
    nb = a.shape[0]
    p = [chain(a.T, b, range(c, nb, 6)) for c in range(6)]
    return ((p[0] + p[2]) + p[4]) + ((p[1] + p[3]) + p[5])


def chain2(a, b):
    """Canonical order of a layer's pre-activation: fused chains over the even and over the odd terms, then their sum."""
    K = a.shape[1]
    return (chain(a, b, range(0, K, 2)) + chain(a, b, range(1, K, 2))).astype(F32)


def forward_canonical(x, We, Wd, alpha):
    zh = chain2(x, We)
    h = lrelu(zh, alpha)
    zy = chain2(h, Wd)
    return zh, h, zy, lrelu(zy, alpha)


def batch_loss(diff, order="blas"):
    """mean over the last axis, then over the rows (App. A.2).  Canonical order: plain sequential float32 sums (row by
    row over the ETFs, then over the rows), as the C restatement writes them."""
    sq = (diff * diff).astype(F32)
    if order == "canonical":
        rows = np.cumsum(sq, axis=1, dtype=F32)[:, -1] / F32(sq.shape[1])
        return F32(np.cumsum(rows, dtype=F32)[-1] / F32(sq.shape[0]))
    return F32(np.mean(np.mean(sq, axis=-1, dtype=F32), dtype=F32))


def mse_and_grads(xb: np.ndarray, We: np.ndarray, Wd: np.ndarray, alpha: float, order: str = "blas"):
    """One train-step forward/backward: loss = sum((y-x)^2)/(B*E) (App. A.2), dWe, dWd."""
    B, E = xb.shape
    canon = order == "canonical"
    zh, h, zy, y = forward_canonical(xb, We, Wd, alpha) if canon else forward(xb, We, Wd, alpha)
    diff = y - xb
    # Keras: mean over the last axis, then SUM_OVER_BATCH_SIZE over the rows.
    loss = batch_loss(diff, order)
    a = F32(alpha)
    if canon:
        dzy = (F32(F32(2.0) / F32(B * E)) * diff).astype(F32)     # the MSE gradient's scalar, rounded once
    else:
        dzy = (F32(2.0) * diff / F32(B * E)).astype(F32)
    dzy = np.where(zy > 0, dzy, a * dzy).astype(F32)      # LeakyReluGrad: features > 0 ? g : a*g
    if canon:
        dWd = class_sum(h, dzy)
        WdT = np.ascontiguousarray(Wd.T)
        dh = chain(dzy, WdT, range(0, E, 2)) + chain(dzy, WdT, range(1, E, 2))
        dzh = np.where(zh > 0, dh, a * dh).astype(F32)
        dWe = class_sum(xb, dzh)
    else:
        dWd = (h.T @ dzy).astype(F32)
        dh = dzy @ Wd.T
        dzh = np.where(zh > 0, dh, a * dh).astype(F32)
        dWe = (xb.T @ dzh).astype(F32)
    return loss, dWe, dWd


class Nadam:
    """Keras 2.7 `Nadam` (optimizer_v2/nadam.py `_prepare_local` + `_resource_apply_dense`),
    restated in float32 (App. A.4).  One global step per mini-batch shared by all variables."""

    def __init__(self, cfg: FitConfig, shapes):
        self.c = cfg
        self.t = 0
        self.m_cache = F32(1.0)
        self.m = [np.zeros(s, F32) for s in shapes]
        self.v = [np.zeros(s, F32) for s in shapes]

    def coefficients(self):
        c = self.c
        b1, b2 = F32(c.beta_1), F32(c.beta_2)
        t = F32(self.t + 1)
        nxt = F32(self.t + 2)
        base = F32(c.decay_base)
        dec = F32(c.schedule_decay)
        # float32 pow = the correctly rounded value (evaluated in float64, then rounded): numpy's own
        # float32 power is SIMD code whose last bit depends on the CPU, and 1 - beta_2^t amplifies it
        # 100x during the first steps.
        def pw(x, y):
            return F32(np.power(np.float64(x), np.float64(y)))
        mu_t = b1 * (F32(1.) - F32(.5) * pw(base, dec * t))
        mu_t1 = b1 * (F32(1.) - F32(.5) * pw(base, dec * nxt))
        ms_new = F32(self.m_cache * mu_t)
        ms_next = F32(ms_new * mu_t1)
        vden = F32(1.) - pw(b2, t)
        return mu_t, mu_t1, ms_new, ms_next, vden

    def step(self, params, grads):
        c = self.c
        mu_t, mu_t1, ms_new, ms_next, vden = self.coefficients()
        b1, b2 = F32(c.beta_1), F32(c.beta_2)
        lr, eps = F32(c.lr), F32(c.epsilon)
        for i, (p, g) in enumerate(zip(params, grads)):
            g_prime = g / (F32(1.) - ms_new)
            self.m[i] = b1 * self.m[i] + (F32(1.) - b1) * g
            m_prime = self.m[i] / (F32(1.) - ms_next)
            self.v[i] = b2 * self.v[i] + (F32(1.) - b2) * (g * g)
            v_prime = self.v[i] / vden
            m_bar = (F32(1.) - mu_t) * g_prime + mu_t1 * m_prime
            p -= lr * m_bar / (np.sqrt(v_prime) + eps)
        self.m_cache = ms_new
        self.t += 1


def split_at(n_rows: int, validation_split: float) -> int:
    """Keras `train_validation_split`: floor(N * (1 - split)) rows train, the tail validates;
    taken before any shuffling (App. A.3)."""
    return int(np.floor(n_rows * (1. - validation_split)))


def _c_backend():
    """The C restatement (oracle/c/fit_oracle.c), or None where it cannot be built (no gcc)."""
    try:
        from oracle import c_oracle
        c_oracle._lib()
        return c_oracle
    except Exception:
        return None


def fit(x, We0, Wd0, cfg: FitConfig | None = None, perm=None, record_grads_at=(), order="canonical", backend="auto"):
    """Model.fit(x, x, ...) as the reference calls it (Autoencoder_encapsulate.py:83-96).

    order  : "canonical" (fixed summation order, see the module docstring) or "blas".
    backend: "numpy", "c", or "auto" — the canonical order runs through the bit-identical C restatement when it is
             available (it is 50x faster than the float64 emulation of fused multiply-adds below; tests/test_oracle.py
             checks the two against each other) unless gradients of several steps are asked for.

    x      : (N, E) training rows (already MinMax-scaled by the caller), any float dtype
    We0/Wd0: initial kernels (E, L) / (L, E)
    perm   : None (no shuffling — the deterministic-order parity setting) or an int array
             (epochs, n_fit) holding the row order of every epoch (Keras shuffle=True draws a
             fresh permutation of the fit rows each epoch; here it is an explicit input).
    Returns dict(We, Wd, loss, val_loss, epochs_run, grads{step: (dWe, dWd)}).
    """
    cfg = cfg or FitConfig()
    x = np.ascontiguousarray(x, dtype=F32)
    record_grads_at = tuple(record_grads_at)
    if order == "canonical" and backend in ("auto", "c") and len(record_grads_at) <= 1:
        c = _c_backend()
        if c is None and backend == "c":
            raise RuntimeError("the C restatement could not be built")
        if c is not None and cfg == dataclasses.replace(FitConfig(), epochs=cfg.epochs, batch_size=cfg.batch_size,
                                                        validation_split=cfg.validation_split, patience=cfg.patience,
                                                        early_stop=cfg.early_stop):
            out = c.fit(x, We0, Wd0, epochs=cfg.epochs, perm=perm, canonical=True, batch=cfg.batch_size,
                        patience=cfg.patience, early_stop=cfg.early_stop, validation_split=cfg.validation_split,
                        grad_step=(record_grads_at[0] if record_grads_at else None))
            out.setdefault("grads", {})
            return out
    We = np.array(We0, dtype=F32, copy=True)
    Wd = np.array(Wd0, dtype=F32, copy=True)
    n_fit = split_at(len(x), cfg.validation_split)
    x_fit, x_val = x[:n_fit], x[n_fit:]
    opt = Nadam(cfg, [We.shape, Wd.shape])
    loss_hist, val_hist, grads = [], [], {}
    best, wait = np.inf, 0
    bs = cfg.batch_size
    epochs_run = 0
    for epoch in range(cfg.epochs):
        row_order = np.arange(n_fit) if perm is None else np.asarray(perm[epoch])
        acc, cnt = F32(0.), 0
        for lo in range(0, n_fit, bs):
            xb = x_fit[row_order[lo:lo + bs]]
            loss, dWe, dWd = mse_and_grads(xb, We, Wd, cfg.alpha, order)
            if opt.t in record_grads_at:
                grads[opt.t] = (dWe.copy(), dWd.copy())
            opt.step([We, Wd], [dWe, dWd])
            acc = F32(acc + loss * F32(len(xb)))
            cnt += len(xb)
        loss_hist.append(F32(acc / F32(cnt)))
        vacc, vcnt = F32(0.), 0
        for lo in range(0, len(x_val), bs):
            xb = x_val[lo:lo + bs]
            y = (forward_canonical if order == "canonical" else forward)(xb, We, Wd, cfg.alpha)[3]
            vl = batch_loss(y - xb, order)
            vacc = F32(vacc + vl * F32(len(xb)))
            vcnt += len(xb)
        val = F32(vacc / F32(vcnt)) if vcnt else F32(np.nan)
        val_hist.append(val)
        epochs_run = epoch + 1
        if cfg.early_stop and vcnt:
            # EarlyStopping.on_epoch_end (App. A.6): strict '<', stop needs epoch > 0.  Without validation rows there is
            # no 'val_loss': Keras warns, leaves `wait` alone and never stops.
            wait += 1
            if val < best:
                best, wait = val, 0
            if wait >= cfg.patience and epoch > 0:
                break
    return dict(We=We, Wd=Wd, loss=np.array(loss_hist, F32), val_loss=np.array(val_hist, F32),
                epochs_run=epochs_run, grads=grads)


def glorot_uniform(seed: int, fan_in: int, fan_out: int) -> np.ndarray:
    """Documented initial-weight stream shared by oracle and product tests: kernel (fan_in,
    fan_out) ~ U(-l, l), l = sqrt(6/(fan_in+fan_out)), drawn in float64 from numpy's Philox
    bit generator keyed by `seed` and rounded to float32.  (Keras' own GlorotUniform uses TF's
    RNG, which cannot be reproduced here; initial weights are therefore an explicit input.)"""
    rng = np.random.Generator(np.random.Philox(key=int(seed)))
    lim = glorot_limit(fan_in, fan_out)
    return rng.uniform(-lim, lim, (fan_in, fan_out)).astype(F32)


def initial_weights(seed: int, E: int, L: int):
    """(We0, Wd0) for one task: encoder stream keyed 2*seed, decoder stream 2*seed+1."""
    return glorot_uniform(2 * seed, E, L), glorot_uniform(2 * seed + 1, L, E)


def make_perm(seed: int, epochs: int, n_fit: int) -> np.ndarray:
    """Explicit per-epoch row orders standing in for Keras' shuffle=True (App. A.3)."""
    rng = np.random.Generator(np.random.Philox(key=int(seed)))
    return np.stack([rng.permutation(n_fit) for _ in range(epochs)]).astype(np.int32)


def device_perm(stream_seed: int, epochs: int, n_fit: int) -> np.ndarray:
    """The CUDA path's device-side epoch shuffle restated (include/hfrep.h, `task_seed`): per epoch the stable argsort of
    fmix32(seed*0x9E3779B1 ^ (epoch+1)*0x85EBCA77 ^ (row+1)*0xC2B2AE3D).  A stand-in for Keras' shuffle=True whose RNG
    cannot be reproduced; both sides of a parity check use this table."""
    M = 0xFFFFFFFF
    out = np.empty((epochs, n_fit), np.int32)
    for e in range(epochs):
        keys = []
        for i in range(n_fit):
            h = ((stream_seed * 0x9E3779B1) & M) ^ (((e + 1) * 0x85EBCA77) & M) ^ (((i + 1) * 0xC2B2AE3D) & M)
            h ^= h >> 16; h = (h * 0x85EBCA6B) & M
            h ^= h >> 13; h = (h * 0xC2B2AE35) & M
            h ^= h >> 16
            keys.append(h)
        out[e] = np.argsort(np.array(keys, np.uint64), kind="stable")
    return out
