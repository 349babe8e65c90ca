import ctypes
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_devices():
    """number of CUDA devices the driver sees (0 without a driver), without importing torch"""
    try:
        cu = ctypes.CDLL("libcuda.so.1")
        n = ctypes.c_int(0)
        if cu.cuInit(0) != 0 or cu.cuDeviceGetCount(ctypes.byref(n)) != 0:
            return 0
        return n.value
    except OSError:
        return 0


def pytest_collection_modifyitems(config, items):
    """tests marked `gpu` are skipped, not failed, on a machine without a CUDA device (the product has
    no CPU path: pyn_init raises there)"""
    if _cuda_devices() > 0:
        return
    skip = pytest.mark.skip(reason="no CUDA device: the library has no CPU path")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
