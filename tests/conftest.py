import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (sm_100a); run with `-m gpu` on the GPU box")


@pytest.fixture(scope="session")
def pkg():
    """The product package (ml-recipe-distributed-pytorch_b200/) under its importable alias."""
    from __graft_entry__ import load_package
    return load_package()


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def load_npz(name):
    import numpy as np
    return dict(np.load(os.path.join(GOLDEN, name)))
