import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_libs():
    """Builds the CPU oracles (test infrastructure).  The verbatim-reference build is only
    possible where /root/reference exists; elsewhere the prebuilt oracle/_ref is used if it
    travelled with the snapshot."""
    from oracle import pyoracle as po
    po.build("port")
    po.build("ref")
    po.build("mref")
    po.build("mdrop")   # (not an oracle: the same wrapper over the product's C++ ORBmatcher, for tests/test_cpp_dropin.py)
    return po


@pytest.fixture(scope="session")
def synth():
    import orb_slam3_mac_b200 as pkg
    return pkg.synth
