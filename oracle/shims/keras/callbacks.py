class EarlyStopping:
    """Bookkeeping itself is in oracle.fitcore.fit (App. A.6); this only carries the arguments."""

    def __init__(self, monitor="val_loss", patience=0, mode="auto", min_delta=0, **k):
        assert monitor == "val_loss" and min_delta == 0
        self.monitor, self.patience, self.mode = monitor, patience, mode
