import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    tree = dict(parent=z["parent"], child=z["child"], bl=z["bl"], root=int(z["root"]))
    mass = dict(off=z["off"], edge=z["edge"], pos=z["pos"], mass=z["mass"])
    return tree, mass, z


def rel_err(got, want):
    return float(np.max(np.abs(got - want) / np.maximum(np.abs(want), 1e-300)))


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle
    oracle.port()  # builds the C restatement on demand
    return oracle
