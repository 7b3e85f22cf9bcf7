import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 and the built libspeechpy_b200.so")


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def lib_built():
    from speaker_verification_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def sp(lib_built):
    """The drop-in speechpy package, GPU-backed."""
    import speaker_verification_b200 as pkg
    pkg.install_dropin()
    import speechpy
    return speechpy
