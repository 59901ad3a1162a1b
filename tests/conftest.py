"""pytest configuration: the ``gpu`` marker and import paths."""

import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for path in (ROOT, os.path.join(ROOT, "ska-sdp-func_b200")):
    if path not in sys.path:
        sys.path.insert(0, path)


def pytest_configure(config):
    config.addinivalue_line(
        "markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    """True if the CUDA driver reports at least one device (no torch import)."""
    import ctypes
    try:
        cuda = ctypes.CDLL("libcuda.so.1")
        count = ctypes.c_int(0)
        return (cuda.cuInit(0) == 0
                and cuda.cuDeviceGetCount(ctypes.byref(count)) == 0
                and count.value > 0)
    except OSError:
        return False


HAVE_GPU = _have_gpu()


def pytest_collection_modifyitems(config, items):
    if HAVE_GPU:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
