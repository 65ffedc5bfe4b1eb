"""configs[4] shape (GAN-generated paths as independent tasks): N paths of (168, 36) in MinMax space, each path its own
training block (columns 0..21 = ETFs), L = 8, fixed epochs.  Every task has its own block -> batch-tile mode of kernel (b).
    python profiles/prof_paths.py [n_paths] [epochs] [warps_per_task]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import hfrep
from hfrep import engine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 100
wpt = int(sys.argv[3]) if len(sys.argv) > 3 else 0
E, L, T = 22, 8, 168
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(7)
paths = torch.clamp(torch.randn((n, T, 36), generator=g, device=dev) * .2 + .61, -.13, 1.26)     # SURVEY §8d cfg 5 stand-in
# helper.factor_hf_split on the device: ETF block of every path (kernel (a), column range 0..21)
starts = torch.arange(n, dtype=torch.int32, device=dev) * T
X = engine.window_gather(paths.view(n * T, 36), starts, T, 0, E)                                  # (n, 168, 22)
assert torch.equal(X, paths[:, :, :E])
We0, Wd0 = engine.initial_weights(1, E, L)
W0 = engine.to_device(np.tile(engine.pack_weights([(We0, Wd0)], E, 2 * E * L), (n, 1)), dev)
task_x = torch.arange(n, dtype=torch.int32, device=dev)
for rep in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fit = engine.fit_batched(X, [L] * n, W0, E=E, T=T, task_x=task_x, max_epochs=epochs, early_stop=0, warps_per_task=wpt)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    flop = n * epochs * (10 * 126 * E * L + 4 * 42 * E * L + 2 * 126 * (L + E) + 3 * 126 * E + 42 * (L + 4 * E) + 28 * E * L * 3)
    print(f"rep {rep}: {n} path-tasks x {epochs} epochs: {ms:.1f} ms -> {n / ms * 1e3:.0f} fits/s at {epochs} epochs "
          f"({n / ms * 1e3 * epochs / 1000:.0f} fits/s per 1000 epochs), {flop / ms / 1e9:.1f} TFLOP/s, status {int(fit.status.abs().sum())}")
