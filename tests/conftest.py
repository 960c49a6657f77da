import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference (oracle/_ref). Built where /root/reference exists; shipped prebuilt to the GPU box."""
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref/libzecale_ref.so not built")
    return R.Ref()


@pytest.fixture(scope="session")
def gpu():
    import zecale_b200 as z
    ctx = z.Context()
    yield ctx
    ctx.close()
