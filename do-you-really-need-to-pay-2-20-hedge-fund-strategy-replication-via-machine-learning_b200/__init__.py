"""hfrep — B200-native (sm_100a) drop-in for the replication hot path of
kaiwenShen/Do-You-Really-Need-to-Pay-2-20-...: ``Autoencoder_encapsulate.AE`` / ``Autoencoder`` and
``helper.*`` with the reference's signatures, over hand-written CUDA kernels behind a C ABI
(include/hfrep.h).  Importing the package loads libhfrep.so and raises if it is not built:
there is no CPU fallback.

    from hfrep import AE, helper                       # short alias of this package
    import sys; sys.path.insert(0, hfrep.DROPIN_PATH)  # `from Autoencoder_encapsulate import AE` as in the notebook
"""
import os as _os

from . import _lib                      # noqa: F401  (fails loudly when the CUDA library is missing)
from . import engine, helper, distributed, generator, h5lite
from .Autoencoder_encapsulate import AE, Autoencoder, config
from .engine import (ReplicationData, SweepPlan, run_mode_r, run_mode_w, run_mode_w_wide, fit_batched, fit_windows, predict_metrics,
                     window_gather, window_index)

DROPIN_PATH = _os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "dropin")
LIB_PATH = _lib.LIB_PATH
__all__ = ["AE", "Autoencoder", "config", "helper", "engine", "distributed", "generator", "h5lite", "ReplicationData", "SweepPlan", "run_mode_r", "run_mode_w", "run_mode_w_wide", "fit_windows",
           "fit_batched", "predict_metrics", "window_gather", "window_index", "DROPIN_PATH", "LIB_PATH"]
