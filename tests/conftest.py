import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_sessionstart(session):
    """A clean checkout has no built artefacts (they are git-ignored): compile the CUDA library and the oracle once, exactly as
    __graft_entry__.build() does, so that the suite does not depend on who ran first.  nvcc cross-compiles without a GPU."""
    import blf_b200  # noqa: F401
    from blf_b200 import capi as c
    from oracle import oracle as O
    if not os.path.exists(c.LIB_PATH):
        c.build()
    O.build()


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def capi():
    import blf_b200  # noqa: F401
    from blf_b200 import capi as c
    c.load()
    return c
