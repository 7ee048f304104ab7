import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# The L2-resident multi-step kernel (csrc/kernels_resident.cuh) takes over `fdtd_b200_step` on lattices below 1 Mi cells, which
# is every lattice small enough for an oracle comparison.  The suite switches it OFF by default so that the streaming kernel
# generations (fused / bulk / vec4 / generic) keep being the thing their tests exercise; tests/test_resident_gpu.py switches it
# back on.  (smoke() and bench.py run with the library's default: on.)
os.environ.setdefault("FDTD_B200_RESIDENT", "0")


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
