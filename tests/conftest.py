import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE_PRESENT = os.path.isdir("/root/reference/cleanrl")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def built_lib():
    """libocppo.so, (re)built if stale -- nvcc cross-compiles without a GPU."""
    import __graft_entry__ as g
    g.build()
    from oc_cleanrl_b200 import _abi
    return _abi.lib()


@pytest.fixture(scope="session")
def oracle_lib():
    import ctypes
    import subprocess
    path = os.path.join(ROOT, "oracle", "_ref", "libocppo_oracle.so")
    if not os.path.exists(path):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle")], check=True)
    return ctypes.CDLL(path)


def load_golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN, name + ".npz"))
