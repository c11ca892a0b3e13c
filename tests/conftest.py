import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    import __graft_entry__ as g
    g.build()


@pytest.fixture(scope="session")
def cbox():
    from montecarlopathtracing_b200 import scenes
    return scenes.load("cbox")


@pytest.fixture(scope="session")
def mis():
    from montecarlopathtracing_b200 import scenes
    return scenes.load("veach_mis")
