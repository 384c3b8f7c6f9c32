import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import pyoracle
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def ref():
    import pyoracle
    if not pyoracle.Ref.available():
        pytest.skip("oracle/_ref (reference compiled from /root/reference) not present")
    return pyoracle.Ref()


@pytest.fixture(scope="session")
def gpu():
    """The product path. Fails loudly (no skip, no fallback) if the extension or the device is missing."""
    import biolib_b200
    ctx = biolib_b200.Context(0)
    yield ctx
    ctx.close()
