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


def fun(x):
    """util/approximate/testing.py:25 demo response."""
    return 3 * x[:, 0] + .5 * np.cos(8 * x[:, 0]) + np.sin(5 * x[:, -1])


def synth(n, q, d, sx=0, sz=1):
    """uniform [0,1)^d, data seed 0, query seed 1."""
    x = np.random.default_rng(sx).random((n, d))
    z = np.random.default_rng(sz).random((q, d))
    return x, z


def load_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    x, z = synth(int(g["n"]), int(g["q"]), int(g["d"]))
    assert x.sum() == float(g["x_sum"]) and z.sum() == float(g["z_sum"]), "numpy RNG stream changed"
    return g, x, z


NN_CASES = ["nn_c1", "nn_k5_d16", "nn_k16_d64", "nn_k32_d16", "nn_k3_d2"]


@pytest.fixture(scope="session")
def oracle():
    from oracle import harness
    harness.lib()
    return harness
