import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "timeout(seconds): pytest-timeout's marker (registered here so that it is known without the plugin, too)")


@pytest.fixture(scope="session")
def built():
    """native pieces are built once per session (nvcc cross-compiles without a GPU)"""
    import __graft_entry__ as g

    g.build()
    return True


@pytest.fixture(scope="session")
def kats():
    import json

    with open(os.path.join(ROOT, "tests", "golden", "reference_kats.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def derived():
    import json

    with open(os.path.join(ROOT, "tests", "golden", "derived_vectors.json")) as f:
        return json.load(f)
