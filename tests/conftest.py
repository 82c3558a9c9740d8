import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    import oracle_lib
    return oracle_lib.Oracle("port")


@pytest.fixture(scope="session")
def ref():
    import oracle_lib
    if not oracle_lib.have_ref():
        pytest.skip("oracle/_ref/libphi_cpu_ref.so not built (needs /root/reference; see oracle/build_ref.sh)")
    return oracle_lib.Oracle("ref")
