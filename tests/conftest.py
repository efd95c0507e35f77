import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def pkg():
    """The product package (its name starts with a digit, hence importlib)."""
    return importlib.import_module("3d_gaussian_splatting_b200")


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle_lib import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def oracle64():
    from oracle.oracle_lib import Oracle
    return Oracle(f64=True)


@pytest.fixture(scope="session")
def rasterizer(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    r = pkg.Rasterizer(0)
    yield r
    r.close()
