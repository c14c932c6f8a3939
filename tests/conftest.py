import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def orc():
    """CPU restatement (oracle/ocw_oracle.hpp); built on demand with g++."""
    from oracle import oracle_py
    oracle_py.build(ref=False)
    return oracle_py.orc()


@pytest.fixture(scope="session")
def ref():
    """UNMODIFIED reference compiled into oracle/_ref/libsmsref.so (built where /root/reference exists,
    shipped prebuilt to the GPU box)."""
    from oracle import oracle_py
    if not oracle_py.have_ref():
        if os.path.isdir("/root/reference/src"):
            oracle_py.build(ref=True)
        else:
            pytest.skip("oracle/_ref/libsmsref.so not built and /root/reference absent")
    return oracle_py.ref()
