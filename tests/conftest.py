import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu through gpurun)")


@pytest.fixture(scope="session")
def lib_built():
    """libmsr.so built in-tree (nvcc cross-compiles without a GPU)."""
    from mllm_sparse_retrieval_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.build()
    return o
