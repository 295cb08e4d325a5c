import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def emu_lib():
    """TEST-ONLY host emulation of the kernel logic (same source as the CUDA library, compiled by g++)."""
    import __graft_entry__ as g

    return g.build_emu()


@pytest.fixture(scope="session")
def cuda_lib():
    import __graft_entry__ as g

    if not os.path.exists(g.LIB):
        g.build_cuda()
    return g.LIB
