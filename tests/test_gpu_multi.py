"""Multi-GPU path on real GPUs: two processes, one per GPU, NCCL.  The sharded job (tasks dealt cyclically over the
ranks, one all-gather of packed result records) must return, on every rank, exactly what one GPU returns for the
whole grid — bit for bit, since every task runs the same kernels on the same inputs wherever it lands.
Skipped on a box with a single GPU (`gpurun --gpus 2` runs it)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
import numpy as np, torch, torch.distributed as dist
import hfrep
from conftest import load_golden
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device(f"cuda:{{rank}}")
dist.init_process_group("nccl", init_method="tcp://127.0.0.1:" + os.environ["PORT"], rank=rank, world_size=world, device_id=dev)
inp = load_golden("ref_inputs")
data = hfrep.ReplicationData(inp["etf"], inp["hfd"], inp["rf"], window=24, device=dev)
Ls, seeds = [21, 13, 8, 5, 2], (123, 124, 125)                  # 15 tasks: a ragged split over 2 ranks
tm = {{}}
tasks, fit, ev = hfrep.run_mode_r(data, Ls, seeds=seeds, max_epochs=40, early_stop=0, sharded=True, timings=tm)
assert len(tasks) == 15 and fit.W.shape[0] == 15 and ev.ante.shape[0] == 15 and tm["gather_ms"] > 0
tasks1, fit1, ev1 = hfrep.run_mode_r(data, Ls, seeds=seeds, max_epochs=40, early_stop=0)       # the whole grid on this GPU
assert tasks == tasks1
for a, b in ((fit.W, fit1.W), (fit.loss, fit1.loss), (fit.val_loss, fit1.val_loss), (fit.epochs_run, fit1.epochs_run),
             (ev.ante, ev1.ante), (ev.post, ev1.post), (ev.turnover, ev1.turnover), (ev.metrics, ev1.metrics),
             (ev.quality, ev1.quality), (ev.mask, ev1.mask), (ev.beta, ev1.beta), (ev.status, ev1.status)):
    assert a.dtype == b.dtype and a.shape == b.shape and torch.equal(a.nan_to_num() if a.is_floating_point() else a,
                                                                     b.nan_to_num() if b.is_floating_point() else b)
# rolling refit, (window, seed) tasks
rng = np.random.default_rng(5)
T, E, F = 70, 22, 5
etf = rng.normal(size=(T, 3)) @ rng.normal(size=(3, E)) * .03 + rng.normal(size=(T, E)) * .01
hfd = etf @ (rng.normal(size=(E, F)) * .1) + rng.normal(size=(T, F)) * .004
kw = dict(train_window=40, latent_dim=3, seeds=(7, 8, 9), ols_window=12, device=dev, max_epochs=15, early_stop=0)
w = hfrep.run_mode_w(etf, hfd, np.full(T, 1e-3), sharded=True, **kw)
w1 = hfrep.run_mode_w(etf, hfd, np.full(T, 1e-3), **kw)
assert torch.equal(w["replica"], w1["replica"]) and torch.equal(w["weights"], w1["weights"]) and torch.equal(w["fit"].W, w1["fit"].W)
# wide rolling refit, windows dealt in runs of consecutive windows: a rank uploads only its stretch of the histories
T, E, F = 300, 64, 40
etf = rng.normal(size=(T, 5)) @ rng.normal(size=(5, E)) * .02 + rng.normal(size=(T, E)) * .01
hfd = etf @ (rng.normal(size=(E, F)) * .1) + rng.normal(size=(T, F)) * .004
kw = dict(train_window=252, latent_dim=6, seeds=(4,), ols_window=24, device=dev, max_epochs=6, early_stop=0)
n_win = T - 252 - 1
W0 = hfrep.engine.pack_weights([hfrep.engine.initial_weights(100 + wi, E, 6) for wi in range(n_win)], E, 2 * E * 6)
mine = hfrep.distributed.shard_blocks(n_win, rank, world)
part = hfrep.run_mode_w_wide(etf, hfd, np.full(T, 1e-3), windows=mine, W0=W0[mine], **kw)
whole = hfrep.run_mode_w_wide(etf, hfd, np.full(T, 1e-3), W0=W0, **kw)
assert part["h2d_bytes"] < whole["h2d_bytes"]
got = hfrep.distributed.gather_results(part["replica"][0], n_win, order="blocks")
assert torch.equal(got, whole["replica"][0])
dist.barrier(); dist.destroy_process_group()
print("ok", rank)
"""


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_sharded_job_equals_single_gpu_job(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [subprocess.Popen([sys.executable, str(script)],
                              env=dict(os.environ, RANK=str(r), WORLD_SIZE="2", PORT=str(port)),
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT) for r in range(2)]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=600)[0].decode())
        except subprocess.TimeoutExpired:
            p.kill()
            outs.append("timeout")
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)
    assert all("ok" in o for o in outs)
