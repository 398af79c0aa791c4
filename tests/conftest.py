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
    """The checker: the reference's own jpeg.c build if present (oracle/_ref), else the port."""
    from oracle import oracle_lib

    return oracle_lib.load()


@pytest.fixture(scope="session")
def decoder():
    import image_codecs_b200 as ic

    if ic.cuda_device_count() < 1:
        pytest.skip("no CUDA device")
    d = ic.BatchDecoder(device=0, profile=True)
    yield d
    d.close()
