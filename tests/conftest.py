import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the in-tree libraries once (no-op when up to date)."""
    from wenoext_b200 import build
    build.build_product()
    build.build_setup()
    build.build_oracle()
    build.build_test_hooks()
    build.build_shim_check()
