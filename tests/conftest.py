import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# The fused calls choose their kernels by call size when nothing forces a path (column-resident kernels up to 6144 columns).
# The tests run small cases, so without this every one of them would exercise the small-call kernels only: pin the
# large-domain path (per-band kernels with scratch strips) for the suite; tests/test_gpu_paths.py covers the column-resident
# kernels and the selection by size explicitly (set_fused_path 1 / 2).  Must be set before the library is loaded.
os.environ.setdefault("RB200_PATH", "strips")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def kd_lw():
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
    return synthetic_kdist("lw")


@pytest.fixture(scope="session")
def kd_sw():
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
    return synthetic_kdist("sw")
