"""OLS(endog, exog).fit().params restated as params = pinv(exog) @ endog, no intercept
(statsmodels' default `method='pinv'`).  PARITY UNPINNED: statsmodels is not installable
here.  The pseudo-inverse is taken in float64 (exog is the float32 latent-factor block of
Autoencoder_encapsulate.py:149; statsmodels would keep float32 and run LAPACK sgesdd, whose
rounding nobody can reproduce — float64 is the mathematically exact value that result
approximates).  `params` is a DataFrame when endog is one, index x1..xL,
columns = endog's columns — helper.py:102 asserts on that type.
"""
import numpy as np
import pandas as pd


class _Result:
    def __init__(self, params):
        self.params = params


class OLS:
    def __init__(self, endog, exog):
        self.endog, self.exog = endog, exog

    def fit(self):
        X = np.asarray(self.exog, dtype=np.float64)
        Y = np.asarray(self.endog, dtype=np.float64)
        beta = np.linalg.pinv(X) @ Y
        if isinstance(self.endog, pd.DataFrame):
            beta = pd.DataFrame(beta, index=[f"x{i + 1}" for i in range(X.shape[1])],
                                columns=self.endog.columns)
        return _Result(beta)
