#!/usr/bin/env python
"""scripts/gpu_fuzz.py on the host emulation of the kernels (tests/emu): the same random parity
cases against the oracle, on the CPU.  TEST TOOL.  usage: python scripts/emu_fuzz.py [n] [first_seed] [wide]"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "scripts"))
from tests.emu_harness import emu_lib  # noqa: E402
from ultrack_b200 import _lib  # noqa: E402

_lib._LIB = emu_lib()
_lib.get_lib = lambda: _lib._LIB
import gpu_fuzz  # noqa: E402

if __name__ == "__main__":
    sys.exit(gpu_fuzz.main())
