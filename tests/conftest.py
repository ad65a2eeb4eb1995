import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def oracle():
    import oracle as _oracle  # oracle/oracle.py -- the checker
    _oracle.lib()
    return _oracle


@pytest.fixture(scope="session")
def b200():
    """The product package (anofox-statistics_b200/), loaded through the repo's own loader."""
    from __graft_entry__ import load_package
    return load_package()
