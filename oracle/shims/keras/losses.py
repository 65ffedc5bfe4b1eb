class MeanSquaredError:
    pass
