import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (runs on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # keep test_zz_* files last whatever order they were collected in
    items.sort(key=lambda it: os.path.basename(str(it.fspath)).startswith("test_zz_"))


@pytest.fixture(scope="session")
def oracle():
    import _oracle
    if not _oracle.available():
        pytest.skip("oracle/_ref/libpim_ref.so not built (needs /root/reference at build time)")
    _oracle.set_seed(1)
    return _oracle
