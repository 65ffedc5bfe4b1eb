"""`from helper import *` exactly as autoencoder_v4.ipynb does: put this directory on sys.path."""
from hfrep.helper import (normalization, read_csv, dic_read, set_seed, random_sampling, transaction_cost,  # noqa: F401
                          price_impact, reshape_cab, ex_post_return, factor_hf_split, dic_save,
                          replica_metrics, series_statistics, data_analysis, annualized_sharpe_ratio, tracking_error, correlation)
