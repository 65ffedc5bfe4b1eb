"""`from Autoencoder_encapsulate import AE` exactly as autoencoder_v4.ipynb does: put this
directory on sys.path instead of the reference checkout."""
import hfrep as _pkg

AE, Autoencoder, config = _pkg.AE, _pkg.Autoencoder, _pkg.config
from hfrep.helper import normalization, price_impact, transaction_cost, reshape_cab, ex_post_return  # noqa: E402,F401
