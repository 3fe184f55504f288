import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HERE = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, HERE):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_mod():
    """The CPU oracle (oracle/): test infrastructure, the checker of every parity test."""
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def gpu():
    """The product package with its CUDA library loaded; fails loudly if the device or library is missing."""
    import ecolab_b200
    ecolab_b200.lib()
    return ecolab_b200

# verbatim driver scripts of the reference (fixtures run by test_gpu_ref_scripts.py), not test modules
collect_ignore = ["ref_scripts"]
