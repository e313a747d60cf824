import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 with `-m gpu`)")


@pytest.fixture(scope="session")
def native_lib():
    """Builds libpdt_b200.so if needed; GPU tests must run on the native path, never a fallback."""
    from pytorch_distributed_template_b200 import build
    build.build()
    from pytorch_distributed_template_b200.ops import _native
    return _native.require()
