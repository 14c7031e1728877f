#!/usr/bin/env python
"""Random link cases on the host emulation of the kernels against oracle.link_oracle (TEST TOOL):
shape, motion, max_distance, max_neighbors, distance_weight, scale, shifts, weight function,
positive-only rule.  usage: python scripts/emu_link_fuzz.py [n] [first_seed]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from tests import cases  # noqa: E402
from tests.emu_harness import emu_lib, segment_frame  # noqa: E402
from ultrack_b200 import synth  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
    first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    lib = emu_lib()
    bad = 0
    for seed in range(first, first + n):
        rng = np.random.default_rng(seed)
        nd = int(rng.integers(2, 4))
        shape = tuple(int(v) for v in (rng.integers(90, 200, 2) if nd == 2 else rng.integers(18, 56, 3)))
        radius = float(rng.uniform(3.5, 6.5))
        tabs = []
        try:
            for t in range(2):
                fg, ct, _ = synth.make_frame(shape, t=t, seed=seed, variant="S", radius=radius, fill=float(rng.uniform(0.25, 0.5)),
                                             motion=float(rng.uniform(0.0, 4.0)))
                tabs.append(segment_frame(lib, fg, ct, keep_labels=True, min_area=int(rng.integers(40, 80)), max_area=4000))
            kw = dict(max_distance=float(rng.uniform(4.0, 18.0)), max_neighbors=int(rng.integers(1, 8)),
                      dw=float(rng.choice([0.0, 0.0, 0.01, 0.1])),
                      scale=None if rng.random() < 0.5 else tuple(float(v) for v in rng.uniform(0.5, 3.0, nd)),
                      weight_func=str(rng.choice(["iou", "iou", "intersection"])), positive=bool(rng.random() < 0.3))
            if rng.random() < 0.4:
                kw["shift"] = rng.normal(0.0, 2.0, (len(tabs[1][1]["area"]), 3)) * (np.arange(3) >= 3 - nd)
            cases.link_and_compare(lib, tabs, nd, min_links=1, max_tie_fraction=0.5, **kw)
        except Exception as e:  # noqa: BLE001
            bad += 1
            print("FAIL", seed, shape, repr(e)[:300], flush=True)
        finally:
            for h, _ in tabs:
                lib.ulk_frame_nodes_free(h)
    print(f"emu link fuzz: {n} cases from seed {first}, {bad} failures")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
