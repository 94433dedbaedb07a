"""pytest configuration: `gpu` marker + import paths"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


def pytest_collection_modifyitems(config, items):
    import torch  # pylint: disable=import-outside-toplevel

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_methods():
    import numpy as np  # pylint: disable=import-outside-toplevel

    return np.load(os.path.join(GOLDEN, "methods.npz"))


@pytest.fixture(params=["oracle", pytest.param("cuda", marks=pytest.mark.gpu)])
def be(request):
    """module with the oracle's call signatures: the C oracle itself, or the CUDA kernels
    behind the C-ABI (numpy in/out adapter)"""
    if request.param == "oracle":
        from oracle import sdm_oracle  # pylint: disable=import-outside-toplevel

        return sdm_oracle
    from . import device_adapter  # pylint: disable=import-outside-toplevel

    return device_adapter
