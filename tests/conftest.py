import os
import sys

# In-process ranks on ONE GPU (tests/test_gpu_slab.py, test_gpu_zzz_inprocess_mailbox.py): the mailbox kernels of one
# rank wait on the device for a kernel of another rank.  With CUDA's default LAZY module loading the first launch
# of a kernel loads its code, which waits for the device to drain -- i.e. for the very kernel that is waiting for it:
# measured 1-2 device-side timeouts per run of the mailbox file with LAZY, none in 6 runs with EAGER
# (scripts/gpu_flaky.sh).  Must be in the environment before CUDA starts.
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

import pytest  # noqa: E402

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
