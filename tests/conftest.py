import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(autouse=True)
def _seeded_lhs_stub():
    """smt is not installed; both sides take the initial swarm from the same seedable stub."""
    from oracle import ref_shim
    ref_shim.install_smt_stub()
    yield


def case(npz, name):
    """Sub-dictionary of the 'name__key' entries of a golden archive."""
    pre = name + "__"
    return {k[len(pre):]: npz[k] for k in npz.files if k.startswith(pre)}
