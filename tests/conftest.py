"""pytest configuration: markers, import paths, build-on-demand of the oracle (gcc) — never of CUDA code."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_ROOT = os.path.join(ROOT, "accv-lab_b200")
for p in (ROOT, PKG_ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure). Built with gcc on first use."""
    from oracle import build_oracle, oracle as orc

    build_oracle.build_c_oracle()
    return orc


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    d = os.path.join(ROOT, "tests", "golden")

    def load(name):
        return dict(np.load(os.path.join(d, name + ".npz")))

    return load


@pytest.fixture(scope="session")
def cabi():
    """ctypes view of libdh_b200.so (the C ABI). Loading needs libcudart only; no GPU is touched until a call."""
    import _cabi

    return _cabi.load()


@pytest.fixture(scope="session")
def cuda_dev():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)
