"""Builds the oracle's C restatement (TEST INFRASTRUCTURE ONLY).

`python -m oracle.build` compiles oracle/ws_oracle.c into oracle/libws_oracle.so
with the system gcc.  The reference (`/root/reference`) is pure Python whose
arithmetic lives in third-party wheels (higra, scikit-image) that are neither
vendored nor installable here, so there is no `oracle/_ref` build: the reference
is unbuildable on this box ("Oracle").
"""
from __future__ import annotations

import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = HERE / "ws_oracle.c"
LIB = HERE / "libws_oracle.so"


def build(force: bool = False) -> Path:
    if LIB.exists() and not force and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    cmd = [
        "gcc", "-O2", "-fPIC", "-shared", "-fvisibility=hidden", "-std=c11",
        "-fno-fast-math", "-ffp-contract=off", "-o", str(LIB), str(SRC), "-lm",
    ]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
