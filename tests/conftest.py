import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def reflib():
    import oracle_libs
    if not oracle_libs.have_ref():
        pytest.skip("oracle/_ref/libdaala_ref.so not built (reference not mounted)")
    return oracle_libs.ref()
