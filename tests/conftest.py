import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as graft  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def O():
    """The CPU oracle (checker)."""
    o = graft.load_oracle()
    assert o.poseidon_selftest() == 0
    return o


@pytest.fixture(scope="session")
def pkg():
    return graft.load_package()


@pytest.fixture(scope="session")
def P2(pkg):
    """The product's plonky2-shaped host API; creating the context fails loudly without a GPU / library."""
    pkg.plonky2.default_context()
    return pkg.plonky2


@pytest.fixture
def rng(request):
    import zlib
    return np.random.default_rng(zlib.crc32(request.node.name.encode()))


P = 0xFFFFFFFF00000001
