import os
import sys

import pytest

# a protocol bug must fail the test run, not hang the GPU box: flag waits give up after 30 s
os.environ.setdefault("NBDIST_B200_TIMEOUT_MS", "30000")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 GPUs (skipped on the 1-GPU box)")


@pytest.fixture(scope="session")
def native_lib():
    """The C-ABI library; building it is part of __graft_entry__.build()."""
    from nbdistributed_b200 import capi

    return capi.load()
