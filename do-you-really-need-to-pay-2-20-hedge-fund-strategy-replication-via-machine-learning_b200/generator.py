"""The reference's saved generator on the GPU: `load_model(path).predict(noise)` as the notebook calls Keras
(autoencoder_v4.ipynb:1277 `generator = load_model('./GAN/trained_generator/MTTS_GAN_GP....h5')`, :1291
`gen_data = generator.predict(np.random.normal(0, 1, (10, 168, 36)), verbose=0)`), without Keras or h5py: the model file
is read by `h5lite`, the forward pass is one fused CUDA kernel (csrc/gen_kernel.cu).

Models taken: what GAN/MTSS_WGAN_GP.py:221-235 (`build_generator`) builds and saves —
LSTM(sigmoid, return_sequences) [-> LeakyReLU] -> LayerNormalization -> LSTM [-> LeakyReLU] -> LayerNormalization -> Dense.
Anything else raises NotImplementedError naming the layer list found.
"""
from __future__ import annotations

import ctypes as C
import json
import os

import numpy as np
import torch

from . import engine, h5lite
from ._lib import lib, check, ptr, stream_ptr


def read_keras_model(path):
    """[(class_name, config, [weights in Keras' order])] of a Keras 2.x HDF5 model file with one Sequential inside."""
    f = h5lite.File(path)
    cfg = json.loads(f.root.attrs["model_config"])
    seq = cfg
    if cfg["class_name"] != "Sequential":                           # Functional wrapper around the Sequential (MTSS_WGAN_GP.py:231-235)
        inner = [l for l in cfg["config"]["layers"] if l["class_name"] == "Sequential"]
        if len(inner) != 1:
            raise NotImplementedError("expected one Sequential model inside the saved model")
        seq = inner[0]
    data = f.root["model_weights"].datasets()
    order = {"LSTM": ("kernel:0", "recurrent_kernel:0", "bias:0"), "LayerNormalization": ("gamma:0", "beta:0"),
             "Dense": ("kernel:0", "bias:0")}
    layers = []
    for l in seq["config"]["layers"]:
        cls, lc = l["class_name"], l["config"]
        ws = []
        for suffix in order.get(cls, ()):
            hits = [k for k in data if ("/" + lc["name"] + "/") in ("/" + k) and k.endswith("/" + suffix)]
            if len(hits) != 1:
                raise h5lite.H5Error(f"weight {lc['name']}/{suffix}: {len(hits)} datasets match")
            ws.append(np.asarray(data[hits[0]], np.float32))
        layers.append((cls, lc, ws))
    return layers


class Generator:
    """`keras.Model` as far as the notebook uses it: predict(x, verbose=0)."""

    def __init__(self, layers, device=None):
        self.layers = layers
        body = [(c, cfg, w) for c, cfg, w in layers if c != "InputLayer"]
        kinds = [c for c, _, _ in body]
        want = None
        for lr1 in (0, 1):
            for lr2 in (0, 1):
                if kinds == ["LSTM"] + ["LeakyReLU"] * lr1 + ["LayerNormalization", "LSTM"] + ["LeakyReLU"] * lr2 + ["LayerNormalization", "Dense"]:
                    want = (lr1, lr2)
        if want is None:
            raise NotImplementedError(f"generator layout {kinds} is not the reference's LSTM generator")
        self.lrelu1, self.lrelu2 = want
        by = lambda c: [(cfg, w) for k, cfg, w in body if k == c]
        (c1, l1), (c2, l2) = by("LSTM")
        for c in (c1, c2):
            if not (c["activation"] == "sigmoid" and c["recurrent_activation"] == "sigmoid" and c["return_sequences"] and c["use_bias"]):
                raise NotImplementedError("LSTM settings other than the reference's (sigmoid activations, return_sequences, bias)")
        (n1c, n1), (n2c, n2) = by("LayerNormalization")
        (dc, d), = by("Dense")
        if dc.get("activation", "linear") != "linear":
            raise NotImplementedError("Dense activation other than linear")
        alphas = [cfg["alpha"] for cfg, _ in by("LeakyReLU")] or [.2]
        self.alpha = float(alphas[0])
        self.ln_eps = float(n1c["epsilon"])
        self.F_in, self.units, self.F_out = l1[0].shape[0], l1[1].shape[0], d[0].shape[1]
        assert l2[0].shape == (self.units, 4 * self.units) and n2c["epsilon"] == n1c["epsilon"]
        packed = np.concatenate([a.ravel() for a in (l1[0], l1[1], l1[2], n1[0], n1[1], l2[0], l2[1], l2[2], n2[0], n2[1], d[0], d[1])])
        assert packed.size == lib.hfrep_generator_weight_count(self.F_in, self.units, self.F_out)
        self._packed_host = packed.astype(np.float32)
        self._device = device
        self._packed_dev = None
        self._tc_dev = None

    def predict_device(self, noise: torch.Tensor, tensor_cores: bool | None = None) -> torch.Tensor:
        """noise f32 [n][T][F_in] on the GPU -> paths f32 [n][T][F_out] on the GPU (no host round trip).
        tensor_cores: True = the tcgen05 kernel (3xTF32 products, FP32 accumulation; the reference's shape only), False = the
        FP32 FFMA kernel, None = tensor cores when the shape allows and there are enough paths to fill the GPU's CTAs."""
        dev = noise.device
        if self._packed_dev is None or self._packed_dev.device != dev:
            self._packed_dev = engine.to_device(self._packed_host, dev)
            self._tc_dev = None
        noise = noise.contiguous().to(torch.float32)
        n, T, Fi = noise.shape
        if Fi != self.F_in:
            raise ValueError(f"expected (n, T, {self.F_in}) noise, got {tuple(noise.shape)}")
        out = torch.empty((n, T, self.F_out), dtype=torch.float32, device=dev)
        tc_bytes = lib.hfrep_generator_tc_weight_bytes(self.F_in, self.units, self.F_out)
        if tensor_cores is None:
            tensor_cores = tc_bytes != 0 and n >= 64 * torch.cuda.get_device_properties(dev).multi_processor_count \
                and os.environ.get("HFREP_GEN_FFMA") is None
        if tensor_cores and tc_bytes == 0:
            raise NotImplementedError("the tensor-core generator kernel takes units = 100, F_in <= 40, F_out <= 36")
        with torch.cuda.device(dev):
            if tensor_cores:
                if self._tc_dev is None:
                    self._tc_dev = torch.empty(tc_bytes, dtype=torch.uint8, device=dev)
                    check(lib.hfrep_generator_tc_prepare(ptr(self._packed_dev), self.F_in, self.units, self.F_out, ptr(self._tc_dev),
                                                         stream_ptr()))
                check(lib.hfrep_generator_forward_tc(ptr(self._packed_dev), ptr(self._tc_dev), self.F_in, self.units, self.F_out,
                                                     self.lrelu1, self.lrelu2, self.alpha, self.ln_eps, ptr(noise), n, T, ptr(out),
                                                     stream_ptr()))
            else:
                check(lib.hfrep_generator_forward(ptr(self._packed_dev), self.F_in, self.units, self.F_out, self.lrelu1, self.lrelu2,
                                                  self.alpha, self.ln_eps, ptr(noise), n, T, ptr(out), stream_ptr()))
        return out

    def predict(self, x, verbose=0, **kwargs):
        dev = engine._dev(self._device)
        with torch.cuda.device(dev):
            out = self.predict_device(engine.to_device(np.asarray(x, np.float32), dev))
            return out.cpu().numpy()

    __call__ = predict


def layers_from_fixture(npz) -> list:
    """Rebuild `read_keras_model`'s result from an .npz written by oracle/make_golden.generator_fixture (layer list as JSON,
    weights as w_<layer>_<index>) — where the model file itself is not available."""
    spec = json.loads(str(npz["layers_json"]))
    layers = []
    for i, (cls, cfg) in enumerate(spec):
        ws, j = [], 0
        while f"w_{i}_{j}" in npz:
            ws.append(np.asarray(npz[f"w_{i}_{j}"], np.float32))
            j += 1
        layers.append((cls, cfg, ws))
    return layers


def load_model(path, device=None) -> Generator:
    """keras.models.load_model for the reference's saved generators (autoencoder_v4.ipynb:1277)."""
    return Generator(read_keras_model(path), device=device)
