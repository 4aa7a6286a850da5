import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG_NAME = "aws-k8s-ansible-provisioner_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with `-m gpu` on the GPU box")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module(PKG_NAME)


@pytest.fixture(scope="session")
def lib(pkg):
    """The built C-ABI library.  Built on demand here (nvcc cross-compiles without a GPU)."""
    if not os.path.exists(pkg.LIB_PATH):
        importlib.import_module(PKG_NAME + ".build").build()
    return pkg.load_library()
