import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    # the CUDA library is built in-tree; build it if this checkout has none yet (nvcc cross-compiles without a GPU)
    import __graft_entry__ as ge
    if not os.path.isfile(ge.LIB_PATH):
        ge.build()


@pytest.fixture(scope="session")
def ref_inputs():
    d = np.load(os.path.join(GOLDEN, "ref_inputs.npz"))
    return {k: d[k] for k in d.files}


def load_golden(name):
    d = np.load(os.path.join(GOLDEN, name + ".npz"))
    return {k: d[k] for k in d.files}


def rel_err(a, b):
    """max |a-b| / max |b|  — array-level relative error used by every tolerance check."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))
