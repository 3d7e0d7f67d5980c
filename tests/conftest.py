import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    path = os.path.join(ROOT, "tests", "golden", "ntxent_golden.npz")
    data = np.load(path)
    cases = {}
    for key in data.files:
        name, field = key.split("/")
        cases.setdefault(name, {})[field] = data[key]
    return cases


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


def seeded_z(M, D, seed=1234, dtype=np.float32):
    """The benchmark's input: i.i.d. N(0,1), generated once on the host."""
    rng = np.random.default_rng(seed)
    return rng.standard_normal((M, D)).astype(dtype)
