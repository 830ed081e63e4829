import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
PKG = os.path.join(ROOT, "relayrl-prototype_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    so = os.path.join(PKG, "relayrl_b200", "librelayrl_b200.so")
    if not os.path.exists(so):  # nvcc cross-compiles sm_100a without a GPU
        import subprocess
        subprocess.check_call(["make", "-s", "-j8", "-C", os.path.join(PKG, "csrc")])


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def load_golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN, name + ".npz"))
