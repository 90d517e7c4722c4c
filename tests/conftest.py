import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def hkm():
    import hkm_b200  # noqa: F401  (alias of the hyphenated package directory)
    return hkm_b200


@pytest.fixture(scope="session")
def oracle():
    from oracle import hkm_oracle
    hkm_oracle.lib()
    return hkm_oracle
