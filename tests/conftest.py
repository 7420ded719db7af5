import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(params=["oracle", pytest.param("device", marks=pytest.mark.gpu)])
def backend(request):
    """The reference's test cases run against both drivers: the CPU oracle here, the CUDA
    path on the GPU box (tests/backends.py)."""
    import backends
    return backends.make(request.param)
