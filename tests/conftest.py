import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def sdrb():
    """The product binding (ctypes over libsdr_b200.so)."""
    return importlib.import_module("sdr-for-android-lib_b200")


@pytest.fixture(scope="session")
def gpu_ctx(sdrb):
    ctx = sdrb.Context(0)
    yield ctx
    ctx.close()
