import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """CHECKER: oracle/liboracle.so (built by oracle/Makefile / __graft_entry__.build())."""
    import subprocess
    from oracle_lib import ORACLE_SO, Oracle
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", os.path.dirname(ORACLE_SO), "oracle"])
    return Oracle()


@pytest.fixture(scope="session")
def F():
    """PRODUCT: fluidsolver_b200 (libfsgpu.so through its C ABI)."""
    import fluidsolver_b200
    fluidsolver_b200.load()
    return fluidsolver_b200
