import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def zctx():
    """One zkb context on cuda:0 for the whole GPU session; fails loudly without the extension."""
    import zklaim_b200 as zk
    ctx = zk.Context(0)
    yield ctx
    ctx.close()


@pytest.fixture(scope="session")
def ref():
    """The reference-backed oracle (oracle/_ref/libzkref.so, prebuilt by oracle/Makefile)."""
    from oracle import zkref
    if not zkref.available():
        pytest.skip("oracle/_ref/libzkref.so not built (make -C oracle ref, needs /root/reference)")
    zkref.lib()
    return zkref
