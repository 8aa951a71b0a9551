import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# every engine the tests create starts with NaN-filled partial_lh / theta_all / _pattern_lh and 0x7f7f scale_num (lkcuda_create): a pattern
# no kernel covered shows up as NaN instead of passing on whatever the allocation held before
os.environ.setdefault("LKCUDA_POISON", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def cuda_lib():
    """The product library; tests marked gpu fail loudly if it cannot be loaded or no device exists."""
    import ctypes
    import iqtree_b200
    lib = iqtree_b200.load_library()
    n = ctypes.c_int32()
    rc = lib.lkcuda_device_count(ctypes.byref(n))
    assert rc == 0 and n.value > 0, "gpu test without a CUDA device (lkcuda_device_count -> %d, %d devices)" % (rc, n.value)
    return lib


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle.oracle_lib import load_oracle_library
    return load_oracle_library()
