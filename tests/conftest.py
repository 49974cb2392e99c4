import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.pyoracle import Oracle

    return Oracle()


@pytest.fixture(scope="session")
def ref_avx512():
    from oracle import pyoracle

    if not pyoracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return pyoracle.Reference(pyoracle.FLAVOUR_AVX512)


@pytest.fixture(scope="session")
def ref_generic():
    from oracle import pyoracle

    if not pyoracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return pyoracle.Reference(pyoracle.FLAVOUR_GENERIC)
