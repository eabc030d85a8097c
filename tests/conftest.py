import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def built_lib():
    """the product library; built in-tree if missing (nvcc cross-compiles without a GPU)"""
    from owlexabrick_b200 import build
    build.build_lib()
    from owlexabrick_b200 import capi
    return capi


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as o
    o.build()
    return o


GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))
