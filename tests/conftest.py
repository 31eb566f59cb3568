"""pytest configuration: registers the ``gpu`` marker and exposes the product package + the oracle.

``-m "not gpu"`` tests run on a CPU-only box (oracle vs golden vectors, host logic, C-ABI symbol checks);
``-m gpu`` tests are the parity tests proper and call the CUDA kernels through the C ABI.
"""
import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


def load_package():
    """The package directory name has a hyphen, so it is imported by string."""
    return importlib.import_module("xhy-flash-attention_b200")


@pytest.fixture(scope="session")
def fa():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    pkg = load_package()
    importlib.import_module("xhy-flash-attention_b200._capi").load()  # fail loudly if the .so is missing
    return pkg


@pytest.fixture(scope="session")
def capi():
    return importlib.import_module("xhy-flash-attention_b200._capi")
