"""pytest configuration: the `gpu` marker and shared model fixtures."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")
    config.addinivalue_line("markers", "ref: needs the compiled reference under oracle/_ref")


@pytest.fixture(scope="session")
def refcore_mod():
    from oracle import refcore
    if not refcore.available():
        pytest.skip("oracle/_ref not built (run oracle/build_ref.sh where /root/reference exists)")
    return refcore
