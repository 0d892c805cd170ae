import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """the CPU checker (oracle/oracle.py) -- test infrastructure only"""
    import oracle as o
    return o


@pytest.fixture(scope="session")
def nb():
    """the product: ctypes face of libnbody_b200.so"""
    import __graft_entry__ as g
    if not os.path.exists(os.path.join(ROOT, "n-bodysimulation_b200", "libnbody_b200.so")):
        g.build()
    import nbody_b200
    return nbody_b200


@pytest.fixture(scope="session")
def golden():
    d = os.path.join(ROOT, "tests", "golden")
    return {f[:-4]: np.load(os.path.join(d, f)) for f in os.listdir(d) if f.endswith(".npz")}


@pytest.fixture(scope="session")
def cuda(nb):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test on a machine without CUDA; the N-body step has no CPU fallback")
    return torch
