"""Generates tests/golden/*.npz by RUNNING the reference's own numba kernels.

Run in the build container only (`python -m oracle.make_golden`); it reads
/root/reference, which does not exist on the GPU box, so the vectors are
committed.  The reference package cannot be imported (higra etc. are missing),
so the five @njit functions are exec'd from their source text with the
`import higra` line dropped -- the function bodies are executed unmodified:

  ultrack/core/segmentation/vendored/graph.py:9-59   edges_from_mask
  ultrack/core/segmentation/node.py:23-118           intersects, intersection_with_bbox_2d/3d,
                                                     _fast_iou_with_bbox
plus the known-answer vectors of
  ultrack/core/segmentation/_test/test_node.py:8-20  (4 bbox cases).
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

REF = Path("/root/reference/ultrack")
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def _load_reference_kernels():
    ns_graph: dict = {}
    text = (REF / "core/segmentation/vendored/graph.py").read_text()
    text = text.split("def mask_to_graph")[0].replace("import higra as hg\n", "")
    exec(compile(text, "ref_graph", "exec"), ns_graph)
    lines = (REF / "core/segmentation/node.py").read_text().splitlines()
    body = "import numpy as np\nfrom numba import njit, types\n" + "\n".join(lines[22:118])
    ns_node: dict = {}
    exec(compile(body, "ref_node", "exec"), ns_node)
    return ns_graph["edges_from_mask"], ns_node


def main() -> None:
    OUT.mkdir(parents=True, exist_ok=True)
    edges_from_mask, node = _load_reference_kernels()
    rng = np.random.default_rng(1234)

    # ---- edges_from_mask: every dtype class the path can see -------------
    cases = {}
    specs = [
        ("f32_3d", (5, 9, 11), np.float32), ("f64_3d", (4, 7, 6), np.float64),
        ("bool_3d", (6, 8, 8), bool), ("u8_3d", (3, 10, 7), np.uint8),
        ("f32_2d", (1, 24, 31), np.float32), ("bool_2d", (1, 17, 19), bool),
        ("f32_thin", (9, 1, 13), np.float32),
    ]
    for name, shape, dt in specs:
        mask = rng.random(shape) < 0.7
        if dt is bool:
            img = rng.random(shape) < 0.4
        elif dt is np.uint8:
            img = rng.integers(0, 256, size=shape, dtype=np.uint8)
        else:
            img = rng.random(shape).astype(dt)
        src, dst, w = edges_from_mask(mask, img, 0.0)
        cases[name + "_mask"] = mask
        cases[name + "_img"] = img
        cases[name + "_src"] = src
        cases[name + "_dst"] = dst
        cases[name + "_w"] = w
    np.savez_compressed(OUT / "edges_cases.npz", **cases)

    # ---- IoU family ---------------------------------------------------------
    iou_cases = {}
    k = 0
    for ndim in (2, 3):
        for _ in range(40):
            lo1 = rng.integers(0, 12, size=ndim)
            lo2 = lo1 + rng.integers(-9, 10, size=ndim)
            s1 = rng.integers(1, 10, size=ndim)
            s2 = rng.integers(1, 10, size=ndim)
            b1 = np.concatenate([lo1, lo1 + s1]).astype(np.int64)
            b2 = np.concatenate([lo2, lo2 + s2]).astype(np.int64)
            m1 = rng.random(tuple(s1)) < 0.7
            m2 = rng.random(tuple(s2)) < 0.7
            m1.flat[0] = True
            m2.flat[0] = True
            val = node["_fast_iou_with_bbox"](b1, b2, m1, m2)
            hit = bool(node["intersects"](b1, b2))
            iou_cases[f"c{k}_b1"], iou_cases[f"c{k}_b2"] = b1, b2
            iou_cases[f"c{k}_m1"], iou_cases[f"c{k}_m2"] = m1, m2
            iou_cases[f"c{k}_iou"] = np.float64(val)
            iou_cases[f"c{k}_hit"] = np.bool_(hit)
            k += 1
    iou_cases["n"] = np.int64(k)
    # known-answer vectors of the reference's own test (test_node.py:8-20)
    ref_vectors = [
        ((0, 0, 5, 5), (1, 1, 2, 2), True),
        ((1, 1, 2, 2), (0, 0, 5, 5), True),
        ((2, 2, 2, 6, 6, 6), (1, 1, 1, 4, 4, 4), True),
        ((5, 5, 7, 7), (3, 2, 6, 4), False),
    ]
    for b1, b2, sol in ref_vectors:
        assert bool(node["intersects"](np.asarray(b1), np.asarray(b2))) == sol
    np.savez_compressed(OUT / "iou_cases.npz", **iou_cases)
    print("wrote", sorted(p.name for p in OUT.glob("*.npz")))


if __name__ == "__main__":
    main()
