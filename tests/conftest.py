import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as ge  # noqa: E402

ge.load_package()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def have_ref():
    """True when the real reference was compiled here (oracle/_ref); never required by -m gpu tests."""
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_gpu_harness"))
