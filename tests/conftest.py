import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def engine():
    """One cgnn_ctx on cuda:0.  On a GPU box a missing library is an error, never a skip."""
    import __graft_entry__ as ge
    ge.build()
    from cgnn_b200 import Engine
    eng = Engine(0)
    yield eng
    eng.close()
