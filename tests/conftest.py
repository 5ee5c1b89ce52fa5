# -*- coding: utf-8 -*-
import json
import os
import sys

import pytest

_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (_root, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope='session')
def golden():
    with open(os.path.join(_root, 'tests', 'golden', 'golden.json')) as fh:
        return json.load(fh)


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope='session')
def cuda_device():
    if not has_cuda():
        pytest.fail("this test is marked gpu but no CUDA device is visible (no CPU fallback exists)")
    return 0
