"""numpy float32 restatement of the reference's saved generator (GAN/MTSS_WGAN_GP.py:221-235 `build_generator`, the
Keras model file loaded at autoencoder_v4.ipynb:1277 and run at :1291) — TEST INFRASTRUCTURE (see oracle/__init__.py).

PARITY UNPINNED: Keras cannot be installed here, and the noise the notebook fed the generator was not saved, so its stored
output (GAN/generated_data2022-07-09.pkl) cannot be regenerated; Keras 2.7 layer semantics are restated from their
definition and the restatement is sanity-checked against that file's range and moments (tests/test_oracle.py):
  LSTM (implementation 2): z = x_t W + h_{t-1} U + b; gates i, f, c, o; recurrent_activation sigmoid on i, f, o;
      `activation` (sigmoid in this model) on the candidate and on the cell output: c_t = f c_{t-1} + i act(z_c), h_t = o act(c_t)
  LeakyReLU(alpha): x if x > 0 else alpha x
  LayerNormalization(axis=-1, epsilon): (x - mean) * rsqrt(var + eps) * gamma + beta, biased variance
  Dense: x K + b
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def sigmoid(x):
    return (F32(1.) / (F32(1.) + np.exp(-x, dtype=F32))).astype(F32)


def lstm(x, W, U, b, activation=sigmoid):
    """x (n, T, F) -> h (n, T, units), Keras gate order i, f, c, o."""
    n, T, _ = x.shape
    units = U.shape[0]
    h = np.zeros((n, units), F32)
    c = np.zeros((n, units), F32)
    out = np.empty((n, T, units), F32)
    for t in range(T):
        z = (x[:, t] @ W + h @ U + b).astype(F32)
        i, f = sigmoid(z[:, :units]), sigmoid(z[:, units:2 * units])
        cand, o = activation(z[:, 2 * units:3 * units]), sigmoid(z[:, 3 * units:])
        c = (f * c + i * cand).astype(F32)
        h = (o * activation(c)).astype(F32)
        out[:, t] = h
    return out


def layer_norm(x, gamma, beta, eps):
    mean = x.mean(axis=-1, keepdims=True, dtype=F32)
    var = ((x - mean) ** 2).mean(axis=-1, keepdims=True, dtype=F32)
    return ((x - mean) / np.sqrt(var + F32(eps)) * gamma + beta).astype(F32)


def leaky_relu(x, alpha):
    return np.where(x > 0, x, F32(alpha) * x).astype(F32)


def forward(layers, noise):
    """layers: [(class_name, config dict, [weight arrays])] in model order (as hfrep.generator.read_keras_model returns)."""
    x = np.asarray(noise, F32)
    for cls, cfg, w in layers:
        if cls == "LSTM":
            assert cfg["activation"] == "sigmoid" and cfg["recurrent_activation"] == "sigmoid" and cfg["return_sequences"]
            x = lstm(x, *w)
        elif cls == "LeakyReLU":
            x = leaky_relu(x, cfg["alpha"])
        elif cls == "LayerNormalization":
            x = layer_norm(x, w[0], w[1], cfg["epsilon"])
        elif cls == "Dense":
            x = (x @ w[0] + w[1]).astype(F32)
        elif cls != "InputLayer":
            raise NotImplementedError(cls)
    return x
