"""No-op stand-in for matplotlib (plots are out of scope; Autoencoder_encapsulate.py:11)."""
