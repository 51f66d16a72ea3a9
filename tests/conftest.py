import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    # files named test_zz_* are collected last by name already; keep collection order stable
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)), it.name) if False else 0)
