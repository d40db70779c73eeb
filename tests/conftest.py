import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    lib_path = os.path.join(entry.PKG_DIR, "libaof_b200.so")
    if not os.path.exists(lib_path):
        entry.build()
    return entry.load_package()


@pytest.fixture(scope="session")
def port():
    from oracle import sad
    return sad.Port()


@pytest.fixture(scope="session")
def ref():
    from oracle import sad
    if not sad.ref_available():
        pytest.skip("oracle/_ref/libaof_ref.so not built (reference tree absent)")
    return sad.Ref()
