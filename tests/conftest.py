import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle.port()


@pytest.fixture(scope="session")
def ref_lib():
    """The compiled reference (oracle/_ref), or None where it was never built."""
    from oracle import pyoracle
    return pyoracle.ref()


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_vectors.npz"))


@pytest.fixture(scope="session")
def cu():
    """The CUDA path under test (C ABI via ctypes); GPU tests only."""
    import torch
    assert torch.cuda.is_available()
    from mllm_b200 import build, capi
    build.build_kernels()
    capi.lib()
    return capi
