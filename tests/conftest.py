import os
import sys

import pytest

# the cross-GPU rendezvous gives up after 10 s by default; the timeout test wants that quickly (read once, at first launch)
os.environ.setdefault("PTB200_BARRIER_TIMEOUT_MS", "300")
# single-device emulation of several ranks = several contexts x (render, merge, copy) streams: give every stream its own
# hardware queue, so that a rendezvous kernel waiting for a peer never sits in front of that peer's kernels (default: 8)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
# ... and load every kernel when the library is loaded: a lazily loaded kernel synchronises the context on its first
# launch, which deadlocks (until the rendezvous times out) when another "rank" of the same process is spinning meanwhile
os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _cuda_available() -> bool:
    try:
        from pathtracing_b200 import _lib
        return _lib.load().ptb_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _cuda_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import pt_oracle
    pt_oracle.lib()
    return pt_oracle
