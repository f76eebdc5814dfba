import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    """A GPU test that hangs (a host-side deadlock, a kernel waiting for a peer) must end the run with a message instead of
    sitting on the box until the outer limit: 25 minutes per test, enforced from a watchdog thread (a signal would not
    get through a blocked C call).  The slowest test (multi-chunk sets with their oracle shards) takes about 3 minutes."""
    if not config.pluginmanager.hasplugin("timeout"):
        return
    for item in items:
        if item.get_closest_marker("gpu") and not item.get_closest_marker("timeout"):
            item.add_marker(pytest.mark.timeout(1500, method="thread"))


@pytest.fixture(scope="session")
def pkg():
    import _pkg

    return _pkg.load()


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def gpu(pkg):
    """One PengCuda handle on cuda:0 for the whole session; fails loudly when the extension is missing."""
    g = pkg.PengCuda(0)
    yield g
    g.close()
