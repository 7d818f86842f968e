"""Test configuration: registers the `gpu` marker, builds the CPU oracle (test infrastructure) once."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    from oracle import shopt_oracle
    shopt_oracle.lib()
    return shopt_oracle


@pytest.fixture(scope="session")
def stage():
    """The product's host API; loading it requires the built libshopt_b200.so (no fallback)."""
    from shopt_b200 import stage as st
    from shopt_b200 import _lib
    _lib.lib()
    return st


def has_cuda():
    try:
        from shopt_b200 import _lib
        return _lib.lib().shopt_device_count() > 0
    except Exception:
        return False
