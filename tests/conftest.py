import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) device; run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def built_lib():
    """The product library; built in-tree when missing (nvcc cross-compiles without a GPU)."""
    from kspider_b200 import capi
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return capi.lib()


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import pairwise_oracle
    return pairwise_oracle.lib()
