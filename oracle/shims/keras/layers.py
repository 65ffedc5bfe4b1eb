import numpy as np

from oracle import fitcore


class Dense:
    """Dense(units, input_dim=..., use_bias=False): kernel (in, out), GlorotUniform."""
    _seed = [123]

    def __init__(self, units, input_dim=None, use_bias=True, **k):
        assert not use_bias, "the reference only builds bias-free layers (Autoencoder_encapsulate.py:24,28)"
        from . import CONTROL
        self.units, self.input_dim = units, input_dim
        if CONTROL["w0"] is not None:
            self.kernel = np.array(CONTROL["w0"](input_dim, units), dtype=np.float32)
        else:
            lim = fitcore.glorot_limit(input_dim, units)
            rng = np.random.Generator(np.random.Philox(key=Dense._seed[0]))
            Dense._seed[0] += 1
            self.kernel = rng.uniform(-lim, lim, (input_dim, units)).astype(np.float32)


class LeakyReLU:
    def __init__(self, alpha=.3, **k):
        self.alpha = alpha
