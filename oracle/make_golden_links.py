"""Generates tests/golden/link_cases.npz by RUNNING the reference's own
`compute_spatial_neighbors` and `color_filtering_mask` bodies.

Build container only (`python -m oracle.make_golden_links`): it reads /root/reference.  ultrack
cannot be imported (higra, sqlalchemy, ... are missing), so the FUNCTION SOURCES are taken from
  ultrack/core/linking/processing.py:20-104   _compute_features, color_filtering_mask
  ultrack/core/linking/processing.py:172-291  compute_spatial_neighbors
  ultrack/core/segmentation/node.py           Node.IoU / intersection / mask_indices /
                                              intensity_mean / intensity_std / _centroid (method bodies)
with `ast`, compiled UNMODIFIED, and executed in a namespace where only the plumbing they
touch is stubbed: `sqla.create_engine` + `DataFrame.to_sql` (the SQL append at the end of the
function), `LOG`.  numpy, scipy.spatial.KDTree, pandas and the numba IoU kernels are the real
ones.  Inputs: nodes cut out of random label images (2-D and 3-D), with and without shifts,
scale, distance weight, the intersection weight / positive-only flags of the GT matcher, and the
colour gate.  Outputs: the (weight, source_id, target_id) rows the reference returns.
"""
from __future__ import annotations

import ast
import logging
import warnings
from contextlib import nullcontext
from pathlib import Path

import numpy as np
import pandas as pd
import scipy.ndimage as ndi
from scipy.spatial import KDTree

REF = Path("/root/reference/ultrack")
OUT = Path(__file__).resolve().parent.parent / "tests" / "golden"


def _function_sources(path: Path, names, cls: str = None):
    text = path.read_text()
    tree = ast.parse(text)
    body = tree.body
    if cls is not None:
        body = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls).body
    out = {}
    for n in body:
        if isinstance(n, ast.FunctionDef) and n.name in names:
            n.decorator_list = []
            out[n.name] = ast.get_source_segment(text, n)
    return out


def load_reference():
    from oracle.make_golden import _load_reference_kernels
    _, node_ns = _load_reference_kernels()
    # ---- Node: the reference's method bodies on a plain class
    methods = _function_sources(REF / "core/segmentation/node.py",
                                {"IoU", "intersection", "mask_indices", "intensity_mean", "intensity_std", "_centroid"},
                                cls="Node")
    import textwrap
    src = "class Node:\n" + "\n".join(textwrap.indent(textwrap.dedent(m), "    ") for m in methods.values())
    ns = dict(np=np, ArrayLike=object, Sequence=object, Tuple=tuple, Optional=object, **{
        k: node_ns[k] for k in ("_fast_iou_with_bbox", "intersects", "intersection_with_bbox_2d", "intersection_with_bbox_3d")})
    exec(compile(src, "ref_node_methods", "exec"), ns)
    Node = ns["Node"]
    # ---- linking functions
    funcs = _function_sources(REF / "core/linking/processing.py",
                              {"_compute_features", "color_filtering_mask", "compute_spatial_neighbors"})

    class _Conn:
        def begin(self):
            return nullcontext(None)

    class _Sqla:
        @staticmethod
        def create_engine(*a, **k):
            return _Conn()

    class _DF(pd.DataFrame):
        def to_sql(self, *a, **k):      # the SQL append is the only side effect: dropped
            return None

    class _Pd:
        DataFrame = _DF

    lns = dict(np=np, pd=_Pd, KDTree=KDTree, warnings=warnings, nullcontext=nullcontext, sqla=_Sqla,
               LOG=logging.getLogger("ref"), Node=Node, List=list, Sequence=object, Optional=object,
               Callable=object, ArrayLike=object, LinkingConfig=object, fasteners=None)
    for name in ("_compute_features", "color_filtering_mask", "compute_spatial_neighbors"):
        exec(compile(funcs[name], f"ref_{name}", "exec"), lns)
    return Node, lns["compute_spatial_neighbors"]


def nodes_from_labels(Node, labels: np.ndarray, time: int, id_base: int):
    out = []
    for lab, sl in enumerate(ndi.find_objects(labels), start=1):
        if sl is None:
            continue
        n = Node()
        n.mask = labels[sl] == lab
        n.bbox = np.array([s.start for s in sl] + [s.stop for s in sl], dtype=np.int64)
        n.id = id_base + lab
        n.time = time
        n.centroid = n._centroid()
        out.append(n)
    return out


class _Cfg:
    def __init__(self, max_distance, max_neighbors, distance_weight, z_score_threshold):
        self.max_distance, self.max_neighbors = max_distance, max_neighbors
        self.distance_weight, self.z_score_threshold = distance_weight, z_score_threshold


def random_labels(shape, seed, n_seeds):
    rng = np.random.default_rng(seed)
    pts = np.stack([rng.integers(0, s, n_seeds) for s in shape], axis=1)
    marks = np.zeros(shape, np.int32)
    for i, p in enumerate(pts, start=1):
        marks[tuple(p)] = i
    dist, idx = ndi.distance_transform_edt(marks == 0, return_indices=True)
    lab = marks[tuple(idx)]
    lab[dist > 5.5] = 0
    return lab


CASES = [
    # name, shape, max_distance, max_neighbors, distance_weight, scale, shift, weight_func, positive, color (z threshold)
    ("plain2d", (64, 64), 12.0, 3, 0.0, None, False, "IoU", False, None),
    ("dw3d", (20, 40, 40), 10.0, 2, 0.05, None, False, "IoU", False, None),
    ("scale_shift3d", (20, 40, 40), 14.0, 3, 0.0, (2.0, 1.0, 1.0), True, "IoU", False, None),
    ("shortscale2d", (64, 64), 12.0, 3, 0.01, (1.5,), True, "IoU", False, None),
    ("gt_intersection3d", (20, 40, 40), 10.0, 3, 0.5, None, False, "intersection", True, None),
    ("positive2d", (64, 64), 12.0, 4, 0.08, None, False, "IoU", True, None),
    ("color2d", (64, 64), 12.0, 3, 0.0, None, False, "IoU", False, 0.8),
    ("color3d_u16", (20, 40, 40), 10.0, 3, 0.0, None, True, "IoU", False, 1.0),
]


def main() -> None:
    Node, csn = load_reference()
    out = {}
    names = []
    for k, (name, shape, md, mn, dw, scale, with_shift, wf, positive, zthr) in enumerate(CASES):
        rng = np.random.default_rng(100 + k)
        l0 = random_labels(shape, 10 + k, 40)
        l1 = np.roll(random_labels(shape, 10 + k, 40), 1, axis=-1)      # same cells moved by one voxel ...
        l1[rng.random(shape) < 0.03] = 0                                # ... and eroded at random
        src = nodes_from_labels(Node, l0, 0, 1_000_000)
        dst = nodes_from_labels(Node, l1, 1, 2_000_000)
        shift = rng.normal(0, 1.2, size=(len(dst), 3)) if with_shift else np.zeros((len(dst), 3))
        images = []
        if zthr is not None:
            dt = np.uint16 if "u16" in name else np.float32
            for c in range(2):
                f = ndi.gaussian_filter(rng.random((2,) + shape), (0, *([3.0] * len(shape))))
                f = (f - f.min()) / (f.max() - f.min()) + 0.08 * rng.random((2,) + shape)
                images.append((f * 4000).astype(dt) if dt == np.uint16 else f.astype(dt))
        bboxes_before = [n.bbox.copy() for n in dst]
        kw = {}
        if wf == "intersection":
            kw["weight_func"] = lambda tgt, s: tgt.intersection(s)
        df = csn(0, _Cfg(md, mn, dw, zthr if zthr is not None else 5.0), src, dst, shift.copy(),
                 None if scale is None else np.asarray(scale), "links", "sqlite://", images,
                 edge_must_be_positive=positive, **kw)
        rows = np.asarray(df, dtype=np.float64).reshape(-1, 3)
        assert len(rows) > 5, (name, len(rows))
        names.append(name)
        out[f"{name}_l0"], out[f"{name}_l1"] = l0, l1
        out[f"{name}_shift"] = shift
        out[f"{name}_rows"] = rows
        out[f"{name}_params"] = np.array([md, mn, dw, -1.0 if zthr is None else zthr, float(positive),
                                          float(wf == "intersection")])
        out[f"{name}_scale"] = np.zeros(0) if scale is None else np.asarray(scale, np.float64)
        for c, im in enumerate(images):
            out[f"{name}_img{c}"] = im
        out[f"{name}_nimg"] = np.int64(len(images))
        for n, b in zip(dst, bboxes_before):       # the reference moves target bboxes in place
            n.bbox = b
    out["names"] = np.array(names)
    np.savez_compressed(OUT / "link_cases.npz", **out)
    print("wrote link_cases.npz:", {n: len(out[f"{n}_rows"]) for n in names})


if __name__ == "__main__":
    main()
