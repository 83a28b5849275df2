import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # make sure the CUDA library and the CPU checkers exist (cross-compiles without a GPU)
    import __graft_entry__ as G

    G.build()


@pytest.fixture(scope="session")
def H():
    import harness

    return harness


@pytest.fixture(scope="session")
def orc(H):
    return H.Oracle()


@pytest.fixture(scope="session")
def ref(H):
    if not H.ref_available():
        pytest.skip("oracle/_ref/libsvtref.so not built (reference tree absent)")
    return H.Ref()
