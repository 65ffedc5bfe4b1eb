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


def mse_and_grads(xb: np.ndarray, We: np.ndarray, Wd: np.ndarray, alpha: float):
    """One train-step forward/backward: loss = sum((y-x)^2)/(B*E) (App. A.2), dWe, dWd."""
    B, E = xb.shape
    zh, h, zy, y = forward(xb, We, Wd, alpha)
    diff = y - xb
    # Keras: mean over the last axis, then SUM_OVER_BATCH_SIZE over the rows.
    loss = F32(np.mean(np.mean(diff * diff, axis=-1, dtype=F32), dtype=F32))
    a = F32(alpha)
    dzy = (F32(2.0) * diff / F32(B * E)).astype(F32)
    dzy = np.where(zy > 0, dzy, a * dzy).astype(F32)      # LeakyReluGrad: features > 0 ? g : a*g
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


def fit(x, We0, Wd0, cfg: FitConfig | None = None, perm=None, record_grads_at=()):
    """Model.fit(x, x, ...) as the reference calls it (Autoencoder_encapsulate.py:83-96).

    x      : (N, E) training rows (already MinMax-scaled by the caller), any float dtype
    We0/Wd0: initial kernels (E, L) / (L, E)
    perm   : None (no shuffling — the deterministic-order parity setting) or an int array
             (epochs, n_fit) holding the row order of every epoch (Keras shuffle=True draws a
             fresh permutation of the fit rows each epoch; here it is an explicit input).
    Returns dict(We, Wd, loss, val_loss, epochs_run, grads{step: (dWe, dWd)}).
    """
    cfg = cfg or FitConfig()
    x = np.ascontiguousarray(x, dtype=F32)
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
        order = np.arange(n_fit) if perm is None else np.asarray(perm[epoch])
        acc, cnt = F32(0.), 0
        for lo in range(0, n_fit, bs):
            xb = x_fit[order[lo:lo + bs]]
            loss, dWe, dWd = mse_and_grads(xb, We, Wd, cfg.alpha)
            if opt.t in record_grads_at:
                grads[opt.t] = (dWe.copy(), dWd.copy())
            opt.step([We, Wd], [dWe, dWd])
            acc = F32(acc + loss * F32(len(xb)))
            cnt += len(xb)
        loss_hist.append(F32(acc / F32(cnt)))
        vacc, vcnt = F32(0.), 0
        for lo in range(0, len(x_val), bs):
            xb = x_val[lo:lo + bs]
            y = forward(xb, We, Wd, cfg.alpha)[3]
            d = y - xb
            vl = F32(np.mean(np.mean(d * d, axis=-1, dtype=F32), dtype=F32))
            vacc = F32(vacc + vl * F32(len(xb)))
            vcnt += len(xb)
        val = F32(vacc / F32(vcnt)) if vcnt else F32(np.nan)
        val_hist.append(val)
        epochs_run = epoch + 1
        if cfg.early_stop:
            # EarlyStopping.on_epoch_end (App. A.6): strict '<', stop needs epoch > 0.
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
