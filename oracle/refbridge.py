"""Import the reference's own `Autoencoder_encapsulate.py` and `helper.py`, unmodified, from
/root/reference over the shim packages in oracle/shims.  TEST INFRASTRUCTURE; runs only in
the build container (the GPU box has no /root/reference) — used by oracle/make_golden.py and
by the CPU tests that re-validate the portable oracle when the reference is present.

Two compat patches, applied to the imported modules / pandas at run time only:
  A. `mean_squared_error(..., squared=False)` (Autoencoder_encapsulate.py:113,130) — the
     kwarg was removed in scikit-learn 1.6; wrap it.
  B. `Series[int]` on a non-integer index (Autoencoder_encapsulate.py:219, helper.py:119,129)
     fell back to positional access before pandas 3; restore that fallback.
"""
from __future__ import annotations

import contextlib
import importlib
import os
import sys

import numpy as np
import pandas as pd

REFERENCE_ROOT = os.environ.get("HFREP_REFERENCE_ROOT", "/root/reference")
_SHIMS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shims")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "Autoencoder_encapsulate.py"))


def _rmse_compat(y_true, y_pred, squared=True, **k):
    from sklearn.metrics import mean_squared_error as mse
    if squared:
        return mse(y_true, y_pred, **k)
    # sklearn < 1.6 with squared=False: sqrt per output column, then the uniform average
    raw = mse(y_true, y_pred, multioutput="raw_values", **k)
    return float(np.mean(np.sqrt(raw)))


@contextlib.contextmanager
def positional_series_getitem():
    """Compat patch B, scoped."""
    orig = pd.Series.__getitem__

    def getitem(self, key):
        if isinstance(key, (int, np.integer)) and not isinstance(key, bool) \
                and not pd.api.types.is_integer_dtype(self.index.dtype):
            return self.iloc[key]
        return orig(self, key)

    pd.Series.__getitem__ = getitem
    try:
        yield
    finally:
        pd.Series.__getitem__ = orig


def load():
    """Returns (Autoencoder_encapsulate module, helper module, keras shim module)."""
    if not available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (REFERENCE_ROOT, _SHIMS, repo):
        if p in sys.path:
            sys.path.remove(p)
    # order: shims (keras/tensorflow/...), then the reference's own files, then the repo for `oracle`
    sys.path[:0] = [_SHIMS, REFERENCE_ROOT, repo]
    for name in ("helper", "Autoencoder_encapsulate"):
        mod = sys.modules.get(name)
        if mod is not None and not getattr(mod, "__file__", "").startswith(REFERENCE_ROOT):
            del sys.modules[name]          # a drop-in of the same name was imported earlier
    helper = importlib.import_module("helper")
    ae = importlib.import_module("Autoencoder_encapsulate")
    assert ae.__file__.startswith(REFERENCE_ROOT) and helper.__file__.startswith(REFERENCE_ROOT)
    ae.mean_squared_error = _rmse_compat            # compat patch A
    keras = importlib.import_module("keras")
    assert keras.__file__.startswith(_SHIMS)
    return ae, helper, keras


def unload():
    """Remove the reference modules and shims from sys.modules / sys.path again."""
    for name in list(sys.modules):
        root = name.split(".")[0]
        if root in ("helper", "Autoencoder_encapsulate", "keras", "tensorflow", "statsmodels", "matplotlib"):
            f = getattr(sys.modules[name], "__file__", "") or ""
            if f.startswith(REFERENCE_ROOT) or f.startswith(_SHIMS):
                del sys.modules[name]
    for p in (_SHIMS, REFERENCE_ROOT):
        while p in sys.path:
            sys.path.remove(p)


def read_reference_data():
    """The three cleaned CSVs the notebook reads (autoencoder_v4.ipynb:71-76), as DataFrames."""
    def rd(name):
        df = pd.read_csv(os.path.join(REFERENCE_ROOT, "cleaned_data", name))
        df["Date"] = pd.to_datetime(df["Date"])
        return df.set_index("Date")
    return rd("factor_etf_data.csv"), rd("hfd.csv"), rd("rf.csv")
