"""Builds the CUDA library in-tree: ultrack_b200/csrc/libultrack_b200.so (sm_100a).

`python -m ultrack_b200.csrc.build [--emu]`.  `--emu` builds the host-emulation harness
used only by the CPU test tier (tests/emu/libulk_emu.so); the package never loads it.
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
SOURCES = ["ulk_rt.cu", "ulk_seg.cu", "ulk_front.cu", "ulk_link.cu", "ulk_aux.cu", "ulk_static.cu", "ulk_api.cu"]
HEADERS = ["ulk_rt.h", "ulk_common.h", "ulk_voxel.h", "ulk_hier.h", "ulk_seg.h", "ulk_front.h", "../../include/ultrack_b200.h"]
LIB = HERE / "libultrack_b200.so"
EMU_LIB = ROOT / "tests" / "emu" / "libulk_emu.so"


def _stale(target: Path) -> bool:
    if not target.exists():
        return True
    mt = target.stat().st_mtime
    return any((HERE / f).stat().st_mtime > mt for f in SOURCES + HEADERS)


def _obj_stale(obj: Path, src: Path) -> bool:
    if not obj.exists():
        return True
    mt = obj.stat().st_mtime
    return src.stat().st_mtime > mt or any((HERE / h).stat().st_mtime > mt for h in HEADERS)


def build(force: bool = False, verbose: bool = False) -> Path:
    """One nvcc -c per source file, in parallel, then a link step (objects under csrc/build/)."""
    if not force and not _stale(LIB):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    flags = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
             "-fmad=false", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC,-fvisibility=hidden"]
    if verbose:
        flags.insert(0, "-Xptxas=-v")
    objdir = HERE / "build"
    objdir.mkdir(exist_ok=True)
    procs, objs = [], []
    for s in SOURCES:
        obj = objdir / (Path(s).stem + ".o")
        objs.append(str(obj))
        if force or _obj_stale(obj, HERE / s):
            procs.append((s, subprocess.Popen([nvcc] + flags + ["-c", "-o", str(obj), str(HERE / s)])))
    failed = [s for s, p in procs if p.wait() != 0]
    if failed:
        raise subprocess.CalledProcessError(1, f"nvcc -c {failed}")
    subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", str(LIB)] + objs, check=True)
    return LIB


def build_emu(force: bool = False) -> Path:
    EMU_LIB.parent.mkdir(parents=True, exist_ok=True)
    if not force and not _stale(EMU_LIB):
        return EMU_LIB
    cmd = ["g++", "-O2", "-std=c++17", "-DULK_EMU", "-fPIC", "-shared", "-ffp-contract=off",
           "-fvisibility=hidden", "-o", str(EMU_LIB)]
    for s in SOURCES:
        cmd += ["-x", "c++", str(HERE / s)]
    subprocess.run(cmd, check=True)
    return EMU_LIB


if __name__ == "__main__":
    if "--emu" in sys.argv:
        print(build_emu(force=True))
    else:
        print(build(force=True, verbose="-v" in sys.argv))
