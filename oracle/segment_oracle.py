"""CPU oracle of ultrack's per-frame segmentation -- TEST INFRASTRUCTURE ONLY.

Literal restatement of (paths relative to /root/reference/ultrack):
  core/segmentation/processing.py:160-340   _process (ids, rows, overlaps)
  core/segmentation/hierarchy.py:33-90      create_hierarchies
  core/segmentation/vendored/hierarchy.py   Hierarchy / oversegment_components
  core/segmentation/vendored/node.py        bbox + mask of a node
  core/segmentation/node.py:229-233         centroid
The higra / scikit-image calls are restated (oracle/ws_oracle.c and below);
"parity unpinned" against real higra -- see DESIGN.md.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Tuple

import numpy as np
import scipy.ndimage as ndi

from oracle.build import build

_KIND = {"area": 0, "dynamics": 1, "volume": 2}

_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(build()))
        i64, dbl, vp = C.c_int64, C.c_double, C.c_void_p
        _lib.ulk_o_edges_from_mask.restype = i64
        _lib.ulk_o_edges_from_mask.argtypes = [vp, vp, i64, i64, i64, C.c_int, vp, vp, vp, vp, vp]
        _lib.ulk_o_hierarchy.restype = i64
        _lib.ulk_o_hierarchy.argtypes = [i64, i64, vp, vp, vp, C.c_int, dbl, dbl, dbl, C.c_int,
                                         vp, vp, vp, vp, vp]
        _lib.ulk_o_cut_labels.restype = i64
        _lib.ulk_o_cut_labels.argtypes = [i64, i64, vp, vp, vp, dbl, vp]
        _lib.ulk_o_leaf_memberships.restype = i64
        _lib.ulk_o_leaf_memberships.argtypes = [i64, i64, vp, vp, i64, vp, vp]
        _lib.ulk_o_watershed_hierarchy.restype = i64
        _lib.ulk_o_watershed_hierarchy.argtypes = [i64, i64, vp, vp, vp, C.c_int, vp, vp]
        _lib.ulk_o_bpt_canonical.restype = i64
        _lib.ulk_o_bpt_canonical.argtypes = [i64, i64, vp, vp, vp, vp, vp, vp]
        _lib.ulk_o_iou.restype = dbl
        _lib.ulk_o_iou.argtypes = [vp, vp, vp, vp]
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


@dataclass
class SegParams:
    """Mirror of the SegmentationConfig fields _process reads
    (config/segmentationconfig.py:19-57)."""

    threshold: float = 0.5
    min_area: int = 100
    max_area: int = 1_000_000
    min_area_factor: float = 4.0
    min_frontier: float = 0.0
    ws_hierarchy: str = "area"


def fix_shift(contours: np.ndarray) -> int:
    """Fixed-point shift of the exact frontier accumulation: weights stay below
    2^E with E derived from the frame's largest finite |contour| value."""
    c = np.asarray(contours)
    if c.size == 0:
        return 60
    c = np.abs(c.astype(np.float64))
    c = c[np.isfinite(c)]
    cmax = float(c.max()) if c.size else 0.0
    e = 0
    while np.ldexp(1.0, e) <= cmax + 1.0:
        e += 1
    return 62 - e


def edges_from_mask(mask: np.ndarray, image: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """vendored/graph.py:9-80 (mask_to_graph + edges_from_mask)."""
    mask = np.ascontiguousarray(mask, dtype=bool)
    if mask.ndim == 2:
        mask = mask[None]
        image = image[None]
    if image.dtype == np.float16:
        image = image.astype(np.float32)
    f32add = int(image.dtype == np.float32)
    img = np.ascontiguousarray(image, dtype=np.float64)
    D, H, W = mask.shape
    n = int(mask.sum())
    ref = np.empty(mask.size, dtype=np.int64)
    src = np.empty(3 * n + 1, dtype=np.int64)
    dst = np.empty(3 * n + 1, dtype=np.int64)
    w = np.empty(3 * n + 1, dtype=np.float32)
    nv = np.zeros(1, dtype=np.int64)
    m8 = mask.view(np.uint8)
    e = lib().ulk_o_edges_from_mask(_p(m8), _p(img), D, H, W, f32add, _p(ref), _p(src), _p(dst), _p(w), _p(nv))
    return src[:e].copy(), dst[:e].copy(), w[:e].copy()


def _hierarchy(n, src, dst, w, prm: SegParams, shift: int):
    N = max(2 * n - 1, 1)
    par = np.empty(N, dtype=np.int64)
    alt = np.empty(N, dtype=np.float64)
    fr = np.empty(N, dtype=np.float64)
    area = np.empty(N, dtype=np.float64)
    hgt = np.empty(N, dtype=np.float64)
    nn = lib().ulk_o_hierarchy(n, len(src), _p(src), _p(dst), _p(w), _KIND[prm.ws_hierarchy],
                               float(prm.min_area), float(prm.max_area), float(prm.min_frontier),
                               shift, _p(par), _p(alt), _p(fr), _p(area), _p(hgt))
    if nn < 0:
        raise RuntimeError("region graph is not connected")
    return par[:nn], alt[:nn], fr[:nn], area[:nn], hgt[:nn]


def label_components(binary: np.ndarray, prm: SegParams) -> np.ndarray:
    """core/segmentation/hierarchy.py:61-73: ndi.label + remove_small_objects."""
    labels, num = ndi.label(binary, output=np.int32)
    if num > 1:
        max_size = int(prm.min_area / prm.min_area_factor) - 1
        sizes = np.bincount(labels.ravel())
        too_small = sizes <= max_size
        labels[too_small[labels]] = 0
    return labels


def oversegment_components(labels: np.ndarray, boundaries: np.ndarray, max_area: int) -> np.ndarray:
    """vendored/hierarchy.py:327-365."""
    dtype = np.promote_types(np.float32, boundaries.dtype)
    boundaries = boundaries.astype(dtype)
    offset = 1
    new_labels = np.zeros_like(labels)
    for lab, sl in enumerate(ndi.find_objects(labels), start=1):
        if sl is None:
            continue
        image = labels[sl] == lab
        if int(image.sum()) > max_area:
            intensity = boundaries[sl] * image
            src, dst, w = edges_from_mask(image, intensity)
            n = int(image.sum())
            leaf = np.empty(n, dtype=np.int64)
            nn = lib().ulk_o_cut_labels(n, len(src), _p(src), _p(dst), _p(w), float(max_area), _p(leaf))
            if nn < 0:
                raise RuntimeError("component graph is not connected")
            # skimage.segmentation.relabel_sequential(cut, offset): 0 is background
            vals = np.unique(leaf)
            vals = vals[vals != 0]
            cut = np.searchsorted(vals, leaf) + offset
            cut[leaf == 0] = 0
            offset = int(cut.max()) + 1
            new_labels[sl][image] = cut
        else:
            new_labels[sl][image] = offset
            offset += 1
    return new_labels


def _generate_id(index: int, time: int, max_segments: int) -> int:
    """core/segmentation/processing.py:42-59."""
    return index + (time + 1) * max_segments


def segment_frame(foreground: np.ndarray, contours: np.ndarray, prm: SegParams, time: int = 0,
                  max_segments_per_time: int = 1_000_000) -> Dict[str, list]:
    """One frame of core/segmentation/processing.py:160-340 (max_noise == 0).

    Returns {"nodes": [dict], "overlaps": [(node_id, ancestor_id)], "labels": int32 array}.
    """
    foreground = np.asarray(foreground)
    contours = np.asarray(contours)
    ndim = foreground.ndim
    binary = foreground > prm.threshold
    labels = label_components(binary, prm)
    labels = oversegment_components(labels, contours, prm.max_area)
    shift = fix_shift(contours)

    nodes: List[dict] = []
    overlaps: List[Tuple[int, int]] = []
    index = 1
    h = -1
    for lab, sl in enumerate(ndi.find_objects(labels), start=1):
        if sl is None:
            continue
        h += 1
        image = labels[sl] == lab
        intensity = contours[sl] * image
        if intensity.dtype == np.float16:
            intensity = intensity.astype(np.float32)
        if intensity.size < 8:
            raise RuntimeError(f"Region too small. Size of {intensity.size} found.")
        n = int(image.sum())
        src, dst, w = edges_from_mask(image, intensity)
        par, alt, fr, area, hgt = _hierarchy(n, src, dst, w, prm, shift)
        nn = len(par)
        emit = np.zeros(nn, dtype=np.uint8)
        emit[n:] = area[n:] <= prm.max_area
        fallback = not emit.any()
        if fallback:
            emit[nn - 1] = 1
        cap = max(n * 8, 64)
        while True:
            o_node = np.empty(cap, dtype=np.int64)
            o_leaf = np.empty(cap, dtype=np.int64)
            c = lib().ulk_o_leaf_memberships(n, nn, _p(par), _p(emit), cap, _p(o_node), _p(o_leaf))
            if c >= 0:
                break
            cap = -c
        o_node, o_leaf = o_node[:c], o_leaf[:c]
        if nn == 1:  # single vertex: the root is the leaf itself
            o_node = np.zeros(1, dtype=np.int64)
            o_leaf = np.zeros(1, dtype=np.int64)
        coords = np.array(np.nonzero(image))  # (ndim, n) in raster == vertex order
        order = np.argsort(o_node, kind="stable")
        o_node, o_leaf = o_node[order], o_leaf[order]
        starts = np.searchsorted(o_node, np.arange(nn), side="left")
        ends = np.searchsorted(o_node, np.arange(nn), side="right")
        hier_index_map: Dict[int, dict] = {}
        for k in np.nonzero(emit)[0]:
            leaves = o_leaf[starts[k]:ends[k]]
            pts = coords[:, leaves]
            lo = pts.min(axis=1)
            hi = pts.max(axis=1) + 1
            mask = np.zeros(tuple(hi - lo), dtype=bool)
            mask[tuple(pts - lo[:, None])] = True
            origin = np.array([s.start for s in sl])
            bbox = np.concatenate([origin + lo, origin + hi]).astype(np.int64)
            # core/segmentation/node.py:229-233
            centroid = (np.array(np.nonzero(mask)).mean(axis=1) + bbox[:ndim]).round().astype(int)
            if ndim == 2:
                z, (y, x) = 0, centroid
            else:
                z, y, x = centroid
            node = dict(
                id=_generate_id(index, time, max_segments_per_time),
                t_node_id=index, t_hier_id=h + 1, t=time,
                z=int(z), y=int(y), x=int(x),
                area=int(n if fallback else area[k]),
                frontier=float(-1.0 if fallback else fr[k]),
                height=float(-1.0 if fallback else hgt[k]),
                bbox=bbox, mask=mask, centroid=centroid,
                hier_parent_id=-1, h_node_index=int(k),
            )
            hier_index_map[int(k)] = node
            nodes.append(node)
            index += 1
        for k, node in hier_index_map.items():
            p = int(par[k])
            if p != k and p in hier_index_map:
                node["hier_parent_id"] = hier_index_map[p]["id"]
            a = k
            while par[a] != a:
                a = int(par[a])
                if a in hier_index_map:
                    overlaps.append((node["id"], hier_index_map[a]["id"]))
        if index > max_segments_per_time:
            raise ValueError(
                f"Number of segments exceeds upper bound of {max_segments_per_time} per time.")
    return {"nodes": nodes, "overlaps": overlaps, "labels": labels}
