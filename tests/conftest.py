import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built_lib():
    import __graft_entry__ as g
    from generalsparse_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        g.build()
    return _lib.lib()
