"""Profiling driver: one fit+eval pass of the cfg2 sweep x SEEDS seeds, no warm-up repeats.
    python profiles/prof_fit.py [seeds] [epochs] [warps_per_task]
Run plain first, then under ncu (see profiles/README.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hfrep
from bench import synthetic_returns, LATENTS

seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 64
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 100
wpt = int(sys.argv[3]) if len(sys.argv) > 3 else 0
etf, hfd, rf = synthetic_returns()
data = hfrep.ReplicationData(etf, hfd, rf, device="cuda:0")
for rep in range(2):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    tasks, fit, ev = hfrep.run_mode_r(data, LATENTS, seeds=range(1000, 1000 + seeds), max_epochs=epochs, early_stop=0,
                                      warps_per_task=wpt)
    b.record()
    torch.cuda.synchronize()
    print(f"rep {rep}: {len(tasks)} tasks x {epochs} epochs: {a.elapsed_time(b):.2f} ms")
