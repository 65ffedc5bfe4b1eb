"""Stand-in for statsmodels: only OLS(Y, X).fit().params is used (Autoencoder_encapsulate.py:151)."""
