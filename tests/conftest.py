import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return entry.load_package()


@pytest.fixture(scope="session")
def ref():
    """oracle/_ref: the reference's own sources compiled for the host (built by __graft_entry__.build())."""
    r = entry.load_oracle()
    if not r.available():
        pytest.skip("oracle/_ref/libradiance_ref.so not built")
    return r


@pytest.fixture(scope="session")
def plt(pkg):
    return pkg.rd.Platform.GetPlatform(0)
