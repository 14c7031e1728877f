"""Test tiers: `-m "not gpu"` runs here (oracle vs golden vectors, host logic, kernel
logic through the host-emulation harness, C-ABI export check); `-m gpu` runs on a B200
and calls the CUDA library through the C ABI."""
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture()
def emu(monkeypatch):
    """Routes ultrack_b200's host layer to the host-emulation harness (tests/emu) so the
    Python side -- rows, blobs, sharding -- is exercised without a GPU.  Test-only."""
    from tests.emu_harness import emu_lib
    from ultrack_b200 import _lib, frames

    lib = emu_lib()
    monkeypatch.setattr(_lib, "_LIB", lib)
    yield lib
    frames.STORE.clear()


@pytest.fixture()
def cuda_lib():
    from ultrack_b200 import _lib, frames

    lib = _lib.get_lib()
    yield lib
    frames.STORE.clear()
