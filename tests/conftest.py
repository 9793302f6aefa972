import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE = "/root/reference"
HAVE_REFERENCE = os.path.isdir(os.path.join(REFERENCE, "generators"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (only in the build container)")


@pytest.fixture
def ref_backend():
    """Install the CPU restatement of the kernels (oracle/kernels_ref.py) under gan_b200.ops for the test."""
    from gan_b200 import kernels
    from oracle.kernels_ref import RefBackend
    prev = kernels.set_backend(RefBackend())
    yield kernels.backend()
    kernels.set_backend(prev)


needs_reference = pytest.mark.skipif(not HAVE_REFERENCE, reason="/root/reference not mounted")
