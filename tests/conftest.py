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
    """The CPU restatement of the reference (test infrastructure)."""
    import oracle as o

    o.build()
    return o


@pytest.fixture(scope="session")
def b200():
    """The product package with its CUDA library loaded.  No fallback: a missing library is an error."""
    import fdtd_b200
    from fdtd_b200 import build

    if not os.path.exists(build.LIB_PATH):
        build.build_extension()
    fdtd_b200.lib()
    return fdtd_b200


def has_gpu() -> bool:
    import fdtd_b200

    return fdtd_b200.lib().fdtd_b200_device_count() > 0
