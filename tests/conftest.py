"""pytest configuration: the ``gpu`` marker and shared helpers."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    """GPU tests are skipped automatically when no CUDA device is visible and
    the run did not select them with ``-m gpu``."""
    try:
        from qiskit_b200 import _lib

        have_gpu = _lib.device_count() > 0
    except Exception:  # library not built yet, or no driver
        have_gpu = False
    if have_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
