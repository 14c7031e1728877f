#!/usr/bin/env python
"""Random parity cases of the CUDA library against the oracle on a GPU box (TEST TOOL, not a test:
the committed GPU tests are deterministic).  Half of the cases use a row length that is a multiple
of 16 so that float32 / float16 / uint8 contours take the TMA front end; the rest take the list
kernels.  usage: python scripts/gpu_fuzz.py [n_cases] [first_seed] [wide]
("wide": 3-D cases span several front-end bricks in every direction)"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from tests import cases  # noqa: E402
from ultrack_b200 import _lib  # noqa: E402
from ultrack_b200.capi import segment_frame  # noqa: E402


def _seg(lib, fg, ct, **kw):
    return segment_frame(lib, fg, ct, keep_labels=True, **kw)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
    wide = len(sys.argv) > 3 and sys.argv[3] == "wide"
    lib = _lib.get_lib()
    bad = 0
    tma = 0
    for seed in range(first, first + n):
        rng = np.random.default_rng(seed)
        nd = int(rng.integers(2, 4))
        shape = [int(v) for v in (rng.integers(16, 200, 2) if nd == 2 else rng.integers(5, 48, 3))]
        if seed % 2 == 0:
            shape[-1] = max(16, (shape[-1] // 16) * 16)
        if wide and nd == 3:
            # several TMA bricks ([16, 4, 128]) in every direction, ragged last bricks
            shape = [int(rng.integers(14, 40)), int(rng.integers(3, 22)), 16 * int(rng.integers(7, 20))]
        shape = tuple(shape)
        kind = cases.KINDS[int(rng.integers(0, len(cases.KINDS)))]
        fg, ct = cases.blob_frame(shape, seed, kind, fill=float(rng.uniform(0.1, 0.9)))
        hier = ("area", "dynamics", "volume")[int(rng.integers(0, 3))]
        min_area = int(rng.integers(2, 120))
        kw = dict(min_area=min_area, max_area=int(min_area * rng.uniform(1.2, 40)), ws_hierarchy=hier,
                  min_area_factor=float(rng.choice([1.0, 2.0, 4.0, 8.0])),
                  min_frontier=float(rng.choice([0.0, 0.0, 0.1, 0.3, 0.6])))
        tma += int(shape[-1] % 16 == 0 and kind != "f64")
        try:
            cases.run_and_compare(lib, _seg, fg, ct, nd, **kw)
        except Exception as e:  # noqa: BLE001
            bad += 1
            print("FAIL", seed, shape, kind, kw, repr(e)[:300], flush=True)
    print(f"gpu fuzz{' (wide)' if wide else ''}: {n} cases from seed {first}, {tma} through the TMA front end, {bad} failures")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
