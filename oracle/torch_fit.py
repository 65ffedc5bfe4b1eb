"""An INDEPENDENT float32 fit on torch-CPU — TEST INFRASTRUCTURE (see oracle/__init__.py).

oracle/fitcore.py restates Keras' fit by hand (forward, backward, Nadam, batching, early stopping) and the C and CUDA
implementations follow it; this module checks that restatement against code none of us wrote: torch autograd for the
backward pass, `torch.nn.functional.leaky_relu` / `mse_loss` for the layers and the loss, and
`torch.optim.NAdam(lr=1e-3, betas=(.9, .999), eps=1e-7, momentum_decay=4e-3)` for the optimiser (formula-identical to
Keras 2.7's Nadam; site-packages torch/optim/nadam.py).  What stays hand-written here is only what
Keras' `fit` does around them (Autoencoder_encapsulate.py:83-96): the chronological validation
tail, consecutive mini-batches of the (optionally permuted) fit rows, row-weighted epoch losses, EarlyStopping.

torch's matmul sums in its own order, so a run agrees with fitcore to ~1e-6 — except where a LeakyReLU pre-activation
passes within rounding of zero (oracle/fitcore.py, module docstring), which shows up as an isolated jump of a few 1e-5
in one epoch's loss that decays again; tests/test_oracle.py states the bound it checks.
"""
from __future__ import annotations

import numpy as np
import torch

from oracle.fitcore import FitConfig, split_at


def fit(x, We0, Wd0, cfg: FitConfig | None = None, perm=None):
    """Same contract as fitcore.fit: returns dict(We, Wd, loss, val_loss, epochs_run)."""
    cfg = cfg or FitConfig()
    torch.set_num_threads(1)                                  # the reference pins TF to one thread (helper.py:38)
    x = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
    n_fit = split_at(len(x), cfg.validation_split)
    x_fit, x_val = x[:n_fit], x[n_fit:]
    We = torch.tensor(np.asarray(We0, np.float32), requires_grad=True)      # Keras kernel layout: (in, out)
    Wd = torch.tensor(np.asarray(Wd0, np.float32), requires_grad=True)
    opt = torch.optim.NAdam([We, Wd], lr=cfg.lr, betas=(cfg.beta_1, cfg.beta_2), eps=cfg.epsilon,
                            momentum_decay=cfg.schedule_decay)
    act = lambda z: torch.nn.functional.leaky_relu(z, negative_slope=cfg.alpha)
    model = lambda xb: act(act(xb @ We) @ Wd)                 # Autoencoder.call, Autoencoder_encapsulate.py:32-35
    loss_hist, val_hist = [], []
    best, wait, epochs_run = float("inf"), 0, 0
    bs = cfg.batch_size
    for epoch in range(cfg.epochs):
        order = torch.arange(n_fit) if perm is None else torch.from_numpy(np.asarray(perm[epoch], np.int64))
        acc, cnt = 0., 0
        for lo in range(0, n_fit, bs):
            xb = x_fit[order[lo:lo + bs]]
            opt.zero_grad(set_to_none=True)
            loss = torch.nn.functional.mse_loss(model(xb), xb)    # mean over all B*E elements (App. A.2)
            loss.backward()
            opt.step()
            acc += float(loss.detach()) * len(xb)
            cnt += len(xb)
        loss_hist.append(acc / cnt)
        with torch.no_grad():
            vacc, vcnt = 0., 0
            for lo in range(0, len(x_val), bs):
                xb = x_val[lo:lo + bs]
                vacc += float(torch.nn.functional.mse_loss(model(xb), xb)) * len(xb)
                vcnt += len(xb)
        val = vacc / vcnt if vcnt else float("nan")
        val_hist.append(val)
        epochs_run = epoch + 1
        if cfg.early_stop and vcnt:                           # EarlyStopping.on_epoch_end (App. A.6)
            wait += 1
            if np.float32(val) < np.float32(best):
                best, wait = val, 0
            if wait >= cfg.patience and epoch > 0:
                break
    return dict(We=We.detach().numpy().copy(), Wd=Wd.detach().numpy().copy(), loss=np.array(loss_hist, np.float32),
                val_loss=np.array(val_hist, np.float32), epochs_run=epochs_run)
