"""configs[3] (synthetic scale-up: 5 000 days x 500 ETFs x 10 000 funds, 252-day windows, rolling refit), a bounded sample
of its 4 747 windows through the wide kernels.   python profiles/prof_wide.py [n_windows] [L] [epochs] [funds]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import hfrep

n_win = int(sys.argv[1]) if len(sys.argv) > 1 else 296
L = int(sys.argv[2]) if len(sys.argv) > 2 else 8
epochs = int(sys.argv[3]) if len(sys.argv) > 3 else 100
F = int(sys.argv[4]) if len(sys.argv) > 4 else 10000
T, E, W = 5000, 500, 252
rng = np.random.default_rng(20240611)                     # SURVEY §8d cfg 4 factor model
fac = rng.normal(size=(T, 10)) * .01
etf = (fac @ rng.normal(size=(10, E)) + rng.normal(size=(T, E)) * .005)
hfd = np.zeros((T, F))
idx = rng.integers(0, E, size=(F, 5))
wts = rng.uniform(.06, .2, size=(F, 5))
for j in range(5):
    hfd += etf[:, idx[:, j]] * wts[:, j]
hfd += rng.normal(size=(T, F)) * .003
rf = np.full(T, 1e-4)
windows = np.linspace(0, T - W - 2, n_win).astype(int)
for rep in range(2):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = hfrep.run_mode_w_wide(etf, hfd, rf, train_window=W, latent_dim=L, seeds=(1,), windows=windows, device="cuda:0",
                                max_epochs=epochs, early_stop=0, timings=(tm := {}))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    ok = int(out["fit"].status.abs().sum()) == 0 and bool(torch.isfinite(out["replica"]).all())
    flop = n_win * epochs * (10 * 189 * E * L + 4 * 63 * E * L)
    print(f"rep {rep}: {n_win} windows x {F} funds, L={L}, {epochs} epochs: {dt * 1e3:.0f} ms end to end (incl. {etf.nbytes + hfd.nbytes >> 20} MiB H2D) "
          f"-> {n_win / dt:.0f} fits/s, {n_win * F / dt / 1e6:.2f} M triples/s, >= {flop / dt / 1e12:.2f} TFLOP/s; ok={ok}; "
          f"fit {tm['fit_ms']:.1f} ms ({flop / tm['fit_ms'] / 1e9:.2f} TFLOP/s), eval {tm['eval_ms']:.1f} ms")
