class Nadam:
    """Defaults only (Autoencoder_encapsulate.py:80); the update is oracle.fitcore.Nadam."""

    def __init__(self, *a, **k):
        assert not a and not k, "the reference calls Nadam() with defaults"
