"""Mode W (rolling refit, north_star's per-(window, seed) task) at the reference's shape: 337 months x 22 ETFs x 13 funds,
168-month training windows -> 168 windows per seed, L = 8.   python profiles/prof_mode_w.py [seeds] [epochs]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hfrep
from bench import synthetic_returns

seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 8
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
etf, hfd, rf = synthetic_returns()
for rep in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = hfrep.run_mode_w(etf, hfd, rf, train_window=168, latent_dim=8, seeds=range(seeds), device="cuda:0",
                           max_epochs=epochs, early_stop=0)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    n = len(out["tasks"])
    print(f"rep {rep}: {n} (window, seed) fits x {epochs} epochs, 13 funds each: {dt * 1e3:.1f} ms end to end -> {n / dt:.0f} fits/s, "
          f"{n * 13 / dt:.0f} (fund, window, seed) triples/s; status {int(out['fit'].status.abs().sum())}")
