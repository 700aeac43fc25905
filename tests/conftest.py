import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ref():
    """The compiled reference (oracle/_ref/libpp_ref.so).  Present in the build container and on
    the GPU box (the .so travels with the snapshot); tests that need it skip otherwise."""
    import refapi
    if not refapi.available():
        pytest.skip("oracle/_ref/libpp_ref.so not built")
    return refapi


@pytest.fixture(scope="session")
def ctx():
    import pathplanning_b200 as pp
    c = pp.Context(0)  # raises loudly without a GPU: there is no CPU fallback
    yield c
    c.close()
