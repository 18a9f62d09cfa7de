import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port():
    from oracle import oracle

    return oracle.port()


@pytest.fixture(scope="session")
def ref():
    """The compiled unmodified reference, or skip when it was never built."""
    from oracle import oracle

    lib = oracle.ref()
    if lib is None:
        pytest.skip("oracle/_ref not built (reference tree not mounted)")
    return lib


@pytest.fixture(scope="session")
def gpu_lib():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from navigation_b200 import library

    return library.load()
