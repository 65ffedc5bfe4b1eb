"""Drop-in for the reference's ``Autoencoder_encapsulate.py``: ``class Autoencoder`` (:19-35) and
``class AE`` (:38-243) keep their names, signatures, defaults, attributes and error behaviour;
the arithmetic runs in the CUDA library (kernel (b) fit, kernel (c) predict + OLS + replica +
metrics, ex-post kernel).  Pandas objects in, pandas objects out, as in the reference.

Differences a caller can see, all inherent to not running TensorFlow:
  * initial kernels and epoch shuffles come from numpy Philox streams restarted by
    ``helper.set_seed`` (Keras used TF's RNG); ``config.shuffle = False`` gives the
    deterministic-order mode used for parity checks;
  * ``train(plot=True)`` prints the summary and plots only if matplotlib is installed.
"""
from __future__ import annotations

import numpy as np
import pandas as pd
import torch

from . import engine, helper
from ._lib import lib, check, ptr, stream_ptr
from .helper import normalization, price_impact, transaction_cost, reshape_cab, ex_post_return  # noqa: F401 (:16)


class config:
    """Knobs the reference hard-codes or leaves to Keras defaults."""
    shuffle = True          # Keras fit(shuffle=True) default; False = deterministic row order
    device = None           # torch device string; None = current CUDA device
    epochs = 1000           # Autoencoder_encapsulate.py:86
    batch_size = 48         # :88
    validation_split = .25  # :89
    beta_index = "first"    # 'first' = the reference (window 0's OLS for every month, :167); 'rolling' = window k for month k


class MinMaxScaler:
    """The slice of sklearn's MinMaxScaler the reference uses (:62-65, :118-119)."""

    def fit(self, x):
        x = np.asarray(x, np.float64)
        self.data_min_, self.data_max_ = x.min(axis=0), x.max(axis=0)
        rng = self.data_max_ - self.data_min_
        rng[rng == 0.] = 1.
        self.scale_ = 1. / rng
        self.min_ = 0. - self.data_min_ * self.scale_
        return self

    def transform(self, x):
        return np.asarray(x, np.float64) * self.scale_ + self.min_

    def fit_transform(self, x):
        return self.fit(x).transform(x)

    def inverse_transform(self, x):
        return (np.asarray(x, np.float64) - self.min_) / self.scale_


class _Half:
    """`Sequential([Dense(no bias), LeakyReLU(.2)])` of :23-30 as seen by callers: predict / get_weights."""

    def __init__(self, owner, which):
        self._owner, self._which = owner, which

    def get_weights(self):
        return [(self._owner.We if self._which == "encoder" else self._owner.Wd).copy()]

    def predict(self, x, verbose=0):
        if self._which == "encoder":
            return self._owner._forward(x, encode_only=True)
        raise NotImplementedError("decoder.predict is not used by the reference; use get_weights()")

    __call__ = predict


class Autoencoder:
    """Autoencoder_encapsulate.py:19-35: E -> latent_dim -> E, no biases, LeakyReLU(.2) after both layers."""

    def __init__(self, latent_dim, input_dim=22):
        self.latent_dim = latent_dim
        self.input_dim = input_dim
        seed = helper.next_stream_seed()
        self._seed = seed
        self.We, self.Wd = engine.initial_weights(seed, input_dim, latent_dim)
        self.encoder = _Half(self, "encoder")
        self.decoder = _Half(self, "decoder")

    def _forward(self, x, encode_only=False):
        x = np.ascontiguousarray(np.asarray(x, np.float32))
        if x.ndim != 2 or x.shape[1] != self.input_dim:
            raise ValueError(f"expected (n, {self.input_dim}) input, got {x.shape}")
        dev = engine._dev(config.device)
        E, L = self.input_dim, self.latent_dim
        with torch.cuda.device(dev):
            W = engine.to_device(engine.pack_weights([(self.We, self.Wd)], E, 2 * E * L), dev)
            xd = engine.to_device(x, dev)
            out = torch.empty((len(x), L if encode_only else E), dtype=torch.float32, device=dev)
            check(lib.hfrep_forward(ptr(W), E, L, ptr(xd), len(x), int(encode_only), .2, ptr(out), stream_ptr()))
            return out.cpu().numpy()

    def call(self, x, **kwargs):
        return self._forward(x)

    __call__ = call

    def predict(self, x, verbose=0):
        return self._forward(x)

    def compile(self, optimizer=None, loss=None, **k):
        pass

    def summary(self):
        E, L = self.input_dim, self.latent_dim
        return (f'Model: "autoencoder"  encoder Dense({E}->{L}, no bias)+LeakyReLU(.2), '
                f'decoder Dense({L}->{E}, no bias)+LeakyReLU(.2); params: {2 * E * L}')


class History:
    def __init__(self, loss, val_loss):
        self.history = {"loss": [float(v) for v in loss], "val_loss": [float(v) for v in val_loss]}
        self.epoch = list(range(len(loss)))


class AE:
    def __init__(self, x_train, y_train, x_test, y_test, latent_dim):
        '''
            data needs to be unscaled.   (signature of Autoencoder_encapsulate.py:39)
        '''
        self.reshape_strat_weight_on_etf = None
        self.window = None
        self.strat_weight_on_etf = None
        self.hfd = None
        self.rf = None
        self.hfd_fullname = None
        self._ante = None
        self._post = None
        self.OOS_hfd = None
        self.OOS_etf = None
        self.OOS_rf = None
        self.test_scale = None
        assert len(x_train) == len(y_train) and len(y_test) == len(x_test)          # :59
        self.history = None
        self.autoencoder = None
        self.train_scale = MinMaxScaler()
        self._x_train = self.train_scale.fit_transform(x_train)                     # :65
        self._x_test = x_test                                                       # :67 (unscaled)
        self._y_train = y_train
        self._y_test = y_test
        self._latent_dim = latent_dim
        self._quality = None
        self._turnover = None

    # -- device staging -----------------------------------------------------------------------
    def _device(self):
        return engine._dev(config.device)

    def _packed(self, dev):
        E, L = self._x_train.shape[1], self._latent_dim
        return engine.to_device(engine.pack_weights([(self.autoencoder.We, self.autoencoder.Wd)], E, 2 * E * L), dev)

    def train(self, patience=5, verbose=2, plot=True):
        '''
        AE training will only use self._x_train        (Autoencoder_encapsulate.py:72-105)
        '''
        E, L = self._x_train.shape[1], self._latent_dim
        self.autoencoder = Autoencoder(L, input_dim=E)
        T = len(self._x_train)
        n_fit = int(np.floor(T * (1. - config.validation_split)))
        dev = self._device()
        with torch.cuda.device(dev):
            X = engine.to_device(self._x_train.astype(np.float32)[None], dev)
            W0 = engine.to_device(engine.pack_weights([(self.autoencoder.We, self.autoencoder.Wd)], E, 2 * E * L), dev)
            # Keras fit(shuffle=True): a fresh row order every epoch, drawn on the device from the model's stream
            shuffle_seeds = [(self.autoencoder._seed + 7) & 0xFFFFFFFF] if config.shuffle else None
            res = engine.fit_batched(X, [L], W0, E=E, T=T, n_fit=n_fit, shuffle_seeds=shuffle_seeds,
                                     max_epochs=config.epochs, batch=config.batch_size, patience=patience, early_stop=1)
            n = int(res.epochs_run[0])
            We, Wd = engine.unpack_weights(res.W[0].cpu().numpy(), E, L)
            loss, val = res.loss[0, :n].cpu().numpy(), res.val_loss[0, :n].cpu().numpy()
        self.autoencoder.We, self.autoencoder.Wd = We, Wd
        self.history = History(loss, val)
        self._quality = None
        if verbose:
            for ep in (range(n) if verbose == 1 else [n - 1]):
                print(f"Epoch {ep + 1}/{config.epochs} - loss: {loss[ep]:.4f} - val_loss: {val[ep]:.4f}")
        if plot:
            print(self.autoencoder.summary())
            try:
                from matplotlib import pyplot as plt
            except ImportError:
                return
            plt.plot(self.history.history['loss'])
            plt.plot(self.history.history['val_loss'])
            plt.title('Model Loss')
            plt.ylabel('accuracy')
            plt.xlabel('epoch')
            plt.legend(['train', 'val'], loc='upper left')
            plt.show()

    # -- kernel (c) ---------------------------------------------------------------------------
    def _evaluate(self, rf_test=None, window=24, want_oos=True, want_weights=False, beta_rolling=False):
        dev = self._device()
        x_test = np.asarray(self._x_test, np.float64)
        y_test = np.asarray(self._y_test, np.float64)
        if y_test.ndim == 1:
            y_test = y_test[:, None]
        if rf_test is None:
            rf_test = np.zeros(len(x_test))
        with torch.cuda.device(dev):
            ev = engine.predict_metrics(
                engine.to_device(self._x_train, dev), engine.to_device(x_test, dev), engine.to_device(y_test, dev),
                engine.to_device(np.asarray(rf_test, np.float64), dev), None, [self._latent_dim], self._packed(dev),
                window=window, want_oos=want_oos, want_weights=want_weights, beta_rolling=beta_rolling)
            torch.cuda.current_stream().synchronize()
        return ev

    def _qual(self):
        if self.autoencoder is None:
            raise AttributeError("'NoneType' object has no attribute 'predict'")   # what :108 would raise
        if self._quality is None:
            window = min(24, len(self._x_test) - 2)
            self._quality = self._evaluate(window=window).quality[0].cpu().numpy()
        return self._quality

    def model_IS_r2(self):          # :107-109
        return float(self._qual()[0])

    def model_IS_RMSE(self):        # :111-113
        return float(self._qual()[1])

    def model_OOS_r2(self):         # :115-122
        q, n_pre = self._qual(), len(self._x_test) - 2
        return [float(v) for v in q[2:2 + n_pre]]

    def model_OOS_RMSE(self):       # :124-131
        q, n_pre = self._qual(), len(self._x_test) - 2
        return [float(v) for v in q[2 + n_pre:2 + 2 * n_pre]]

    def ante(self, rf, hfd, window=24, ):
        '''
        calculate ex-ante return      (Autoencoder_encapsulate.py:133-201)
        '''
        assert isinstance(rf, pd.DataFrame)                                         # :138
        n_test = len(self._x_test)
        n_oos = n_test - window - 1
        rf_tail = np.asarray(rf.iloc[-n_oos:], np.float64).reshape(-1)               # :185
        rf_test = np.zeros(n_test)
        rf_test[n_test - n_oos:] = rf_tail
        ev = self._evaluate(rf_test, window=window, want_oos=False, want_weights=True,
                            beta_rolling=(config.beta_index == "rolling"))
        ante = ev.ante[0].cpu().numpy()
        w = ev.weights[0].cpu().numpy()                                             # (n_oos, E, F)
        self._turnover = ev.turnover[0].cpu().numpy()
        self._replica_metrics = ev.metrics[0].cpu().numpy()
        self.OOS_etf = np.array(self._x_test.iloc[-n_oos:]) if hasattr(self._x_test, "iloc") else np.asarray(self._x_test)[-n_oos:]
        self.OOS_hfd = self._y_test.iloc[-n_oos:] if hasattr(self._y_test, "iloc") else np.asarray(self._y_test)[-n_oos:]
        self.OOS_rf = np.array(rf.iloc[-n_oos:])
        ae_ret_ante = pd.DataFrame(ante)
        ae_ret_ante.columns = hfd.columns                                           # :192
        ae_ret_ante.index = hfd.index[-len(ae_ret_ante):]                           # :193
        self._ante = ae_ret_ante
        self.rf = rf
        self.hfd = hfd
        self.strat_weight_on_etf = [pd.DataFrame(w[k], columns=hfd.columns) for k in range(n_oos)]
        self.reshape_strat_weight_on_etf = [
            pd.DataFrame(w[:, :, f], index=[hfd.columns[f]] * n_oos) for f in range(w.shape[2])]   # reshape_cab, :199
        self.window = window
        return self._ante

    def post(self, factor_etf_data):
        if self._ante is None:
            raise Exception('please execute ante before turnover')                  # :204-205
        OOS_factor_etf = (factor_etf_data.iloc[-len(self.reshape_strat_weight_on_etf[0]) - self.window:])   # :206
        self._post = ex_post_return(self._ante, self.window, self.reshape_strat_weight_on_etf, OOS_factor_etf)
        return self._post

    def turnover(self, hfd_fullname):
        if self._ante is None:
            raise Exception('please execute ante before turnover')                  # :211-212
        turnover = self._turnover                                                   # kernel (c); :213-216
        turnover_df = []
        for i in range(len(turnover)):
            turnover_df.append([list(hfd_fullname.values())[i], turnover[i]])
        turnover_df = pd.DataFrame(turnover_df, columns=['Real_AE', 'Turnover'])
        turnover_df = turnover_df.set_index('Real_AE')
        self.hfd_fullname = hfd_fullname
        return turnover_df

    def plot(self, hfd_fullname, title=None):
        """Cumulative ex-ante / ex-post / real returns per fund (:226-243).  Plotting is outside the
        accelerated path: it needs matplotlib and silently does nothing without it."""
        assert isinstance(title, str)                                               # :227
        try:
            from matplotlib import pyplot as plt
        except ImportError:
            return
        funds = list(self._ante.columns)
        ncol = 3
        nrow = -(-len(funds) // ncol)
        fig, axes = plt.subplots(nrow, ncol, figsize=(30, 4 * nrow), squeeze=False)
        curves = (("Ex-ante", self._ante), ("Ex_post", self._post), ("Real", self.OOS_hfd))
        for j, fund in enumerate(funds):
            a = axes[j // ncol][j % ncol]
            for label, frame in curves:
                a.plot(frame.iloc[:, j].cumsum(), label=label)
            a.legend(loc="upper left")
            a.set_title(hfd_fullname[fund])
        fig.suptitle(title, y=0.93, fontsize=24)
        plt.show()
