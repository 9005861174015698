import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ref_api():
    """The CPU oracle (test infrastructure)."""
    from oracle import loader

    return loader.api()


@pytest.fixture(scope="session")
def gpu_api():
    """The product library; fails loudly if it has not been built."""
    import starframe_b200 as sf

    return sf.product_api()
