"""Profiling driver: the cfg2 sweep (L = 1..21) x SEEDS seeds, kernels (b) and (c) timed separately with
CUDA events, inputs resident.   python profiles/prof_fit.py [seeds] [epochs] [warps_per_task] [reps]
Run plain first, then under ncu (see profiles/README.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hfrep
from hfrep import engine
from bench import synthetic_returns, LATENTS, E, WINDOW

seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 64
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 100
wpt = int(sys.argv[3]) if len(sys.argv) > 3 else 0
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
if os.environ.get("PROF_LATENTS"):
    LATENTS = [int(v) for v in os.environ["PROF_LATENTS"].split(",")]
etf, hfd, rf = synthetic_returns()
data = hfrep.ReplicationData(etf, hfd, rf, device="cuda:0")
tasks = [(L, s) for s in range(1000, 1000 + seeds) for L in LATENTS]
task_L = [L for L, _ in tasks]
W0 = engine.to_device(engine.pack_weights([engine.initial_weights(s, E, L) for L, s in tasks], E, 2 * E * 21), data.device)
best = [float("inf"), float("inf")]
for rep in range(reps):
    a, b, c = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    a.record()
    fit = engine.fit_batched(data.Xtr32, task_L, W0, E=E, T=data.n_train, max_epochs=epochs, early_stop=0, warps_per_task=wpt)
    b.record()
    ev = engine.predict_metrics(data.Xtr64, data.Xte64, data.Yte64, data.rf64, data.sigma, task_L, fit.W, window=WINDOW)
    c.record()
    torch.cuda.synchronize()
    print(f"rep {rep}: {len(tasks)} tasks x {epochs} epochs: fit {a.elapsed_time(b):.3f} ms, eval {b.elapsed_time(c):.3f} ms")
    best = [min(best[0], a.elapsed_time(b)), min(best[1], b.elapsed_time(c))]
print(f"best of {reps}: fit {best[0]:.3f} ms, eval {best[1]:.3f} ms")
