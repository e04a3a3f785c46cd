import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def lbm():
    """The product's Python mirror of the C ABI (lattice-boltzmann_b200/)."""
    return importlib.import_module("lattice-boltzmann_b200")


@pytest.fixture(scope="session")
def oracle():
    """CPU oracle (oracle/lbm_oracle.c): the checker, never the thing measured."""
    import oracle_lib
    oracle_lib.build_oracle()
    return oracle_lib.Oracle(omp=True)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    return load
