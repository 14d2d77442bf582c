import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the native pieces exist (the oracle is tiny; the CUDA library is normally
    prebuilt by __graft_entry__.build() and travels to the GPU box)."""
    from oracle import oracle
    oracle.lib()
    lib = os.path.join(ROOT, "spmm_b200", "libb200sp.so")
    if not os.path.exists(lib):
        import __graft_entry__
        __graft_entry__.build()
