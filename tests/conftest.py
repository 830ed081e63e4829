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


@pytest.fixture(autouse=True)
def _pin_oracle_threads():
    """The checker must be deterministic.  The reference's torch-CPU learner is not reproducible
    across MKL thread counts: on a 24-thread Xeon box the same oracle epoch moved 315 of 17 793 value
    weights by up to 1e-5 relative to its own 8-thread run (while the CUDA path matched the 8-thread
    run to 3e-8 everywhere).  All parity tests therefore run the oracle with 8 threads."""
    try:
        import torch
        torch.set_num_threads(8)
    except Exception:
        pass
    yield
