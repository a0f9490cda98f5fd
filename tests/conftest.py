import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def pytest_collection_modifyitems(config, items):
    # GPU minutes are scarce: no GPU test may hang (pytest-timeout is in the image)
    if config.pluginmanager.hasplugin("timeout"):
        for item in items:
            if item.get_closest_marker("gpu") and not item.get_closest_marker("timeout"):
                item.add_marker(pytest.mark.timeout(420))


@pytest.fixture(scope="session")
def golden():
    from tests.helpers import load_golden
    return load_golden()
