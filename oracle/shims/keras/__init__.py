"""Stand-in for the slice of Keras 2.7 that Autoencoder_encapsulate.py uses (:6-10, :19-35,
:78-96).  The arithmetic lives in oracle/fitcore.py; this file only adapts it to the Keras
object model.  PARITY UNPINNED (see oracle/__init__.py).

CONTROL lets the oracle harness make the run deterministic: `w0(fan_in, fan_out)` supplies
each Dense kernel (default: GlorotUniform from a Philox stream), `perm` the per-epoch row
orders (None = no shuffling), `fit_overrides` FitConfig fields (e.g. epochs).
"""
import numpy as np

from oracle import fitcore

CONTROL = {"w0": None, "perm": None, "fit_overrides": {}, "record_grads_at": (), "last_fit": None}


class Sequential:
    def __init__(self, layers=None):
        self.layers = list(layers or [])

    def _dense(self):
        from .layers import Dense
        return [l for l in self.layers if isinstance(l, Dense)][0]

    def _alpha(self):
        from .layers import LeakyReLU
        return [l for l in self.layers if isinstance(l, LeakyReLU)][0].alpha

    def __call__(self, x):
        return fitcore.lrelu(np.asarray(x, np.float32) @ self._dense().kernel, self._alpha())

    def predict(self, x, verbose=0):
        return self(np.asarray(x, np.float32))

    def get_weights(self):
        return [self._dense().kernel.copy()]


class History:
    def __init__(self):
        self.history = {}


class Model:
    def __init__(self, *a, **k):
        self.optimizer = None
        self.loss = None
        self.stop_training = False

    def compile(self, optimizer=None, loss=None, **k):
        self.optimizer, self.loss = optimizer, loss

    def __call__(self, x):
        return self.call(x)

    def predict(self, x, verbose=0):
        return self.call(np.asarray(x, np.float32))

    def summary(self):
        return "Model: autoencoder (oracle shim)"

    def fit(self, x, y, epochs=1, verbose=0, batch_size=32, validation_split=0., callbacks=()):
        assert x is y or np.array_equal(np.asarray(x), np.asarray(y))
        enc, dec = self.encoder._dense(), self.decoder._dense()
        patience, early = 0, False
        for cb in callbacks:
            patience, early = cb.patience, True
        cfg = fitcore.FitConfig(epochs=epochs, batch_size=batch_size, validation_split=validation_split,
                                patience=patience, early_stop=early, alpha=self.encoder._alpha())
        for k, v in CONTROL["fit_overrides"].items():
            setattr(cfg, k, v)
        out = fitcore.fit(np.asarray(x), enc.kernel, dec.kernel, cfg, perm=CONTROL["perm"],
                          record_grads_at=CONTROL["record_grads_at"])
        enc.kernel, dec.kernel = out["We"], out["Wd"]
        CONTROL["last_fit"] = out
        h = History()
        h.history = {"loss": [float(v) for v in out["loss"]], "val_loss": [float(v) for v in out["val_loss"]]}
        return h
