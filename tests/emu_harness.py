"""Host-emulation harness for the CPU test tier (TEST INFRASTRUCTURE ONLY).

Builds tests/emu/libulk_emu.so from the SAME kernel sources as the CUDA library with
-DULK_EMU (kernel functors executed thread by thread on the host) so that the kernel
logic -- basins, join keys, hierarchy ordering, link selection -- is checked against the
oracle without a GPU.  The package itself never loads this library.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Sequence

import numpy as np

from ultrack_b200 import _lib
from ultrack_b200.csrc.build import build_emu

_EMU = None


def emu_lib():
    global _EMU
    if _EMU is None:
        # ULK_EMU_LIB: a differently built harness (e.g. g++ -fsanitize=address,undefined, run with the
        # sanitizer runtimes in LD_PRELOAD) instead of tests/emu/libulk_emu.so
        _EMU = _lib.bind(C.CDLL(os.environ.get("ULK_EMU_LIB") or str(build_emu())))
        _EMU._ulk_is_emu = True
    return _EMU


from ultrack_b200.capi import link_frames, seg_params, segment_frame  # noqa: E402,F401


from oracle.compare import compare_nodes  # noqa: E402,F401
