"""`properties=` features of `segment()` (create_feats_callback, core/segmentation/processing.py:107-157).

The reference paints every node into a full-frame bool buffer and runs skimage's
`RegionProperties` on it, one node at a time, then concatenates the requested properties into a
float32 vector stored in `nodes.features`.  Here the properties that are reductions over the
node masks are computed for all nodes of a frame at once on the device (`ulk_node_stats`:
per-node intensity mean / std / min / max and the float centroid over the bit-packed masks;
area and bbox come with the node table).  Anything else raises with the supported list.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np

from ultrack_b200 import capi

GEOMETRY = ("area", "num_pixels", "area_bbox", "extent", "bbox", "centroid", "label")
INTENSITY = ("intensity_mean", "intensity_min", "intensity_max", "intensity_std")
SUPPORTED = GEOMETRY + INTENSITY


def _check(properties: Sequence[str], has_image: bool) -> None:
    for p in properties:
        if p not in SUPPORTED:
            raise NotImplementedError(f"property {p!r} is not computed by ultrack_b200; supported: {', '.join(SUPPORTED)}")
        if p in INTENSITY and not has_image:
            raise AttributeError(f"{p!r} needs `images` (skimage: 'No intensity image specified.')")


def n_channels(image, n_frame_dims: int) -> Optional[int]:
    """None: no channel axis (image.ndim == 1 + frame dims); else the size of the last axis."""
    if image is None:
        return None
    return None if len(image.shape) == n_frame_dims + 1 else int(image.shape[-1])


def properties_names(shape: Sequence[int], image, properties: Optional[List[str]]) -> Optional[List[str]]:
    """Column names as `regionprops_table` spells them (_get_properties_names, processing.py:350-387):
    scalars keep their name, vector properties get `-i` suffixes, per-channel intensities `-c`."""
    if properties is None:
        return None
    ndim = len(shape) - 1
    _check(properties, image is not None)
    nch = n_channels(image, ndim)
    names: List[str] = []
    for p in properties:
        if p == "bbox":
            names += [f"bbox-{i}" for i in range(2 * ndim)]
        elif p == "centroid":
            names += [f"centroid-{i}" for i in range(ndim)]
        elif p in INTENSITY and nch is not None:
            names += [f"{p}-{c}" for c in range(nch)]
        else:
            names.append(p)
    return names


def frame_features(lib, handle, nodes: dict, ndim: int, image_frame, properties: List[str]) -> np.ndarray:
    """float32 [n_nodes, n_features]: the rows `_feats_callback` returns, for every node of the frame."""
    _check(properties, image_frame is not None)
    n = len(nodes["area"])
    chans: list = []
    multi = False
    if image_frame is not None and any(p in INTENSITY for p in properties):
        multi = len(image_frame.shape) == ndim + 1
        chans = [image_frame[..., c] for c in range(image_frame.shape[-1])] if multi else [image_frame]
    st = capi.node_stats(lib, handle, chans, centroid="centroid" in properties,
                         minmax=any(p in ("intensity_min", "intensity_max") for p in properties))
    host = {k: v.host() for k, v in st.items()}
    lo = 3 - ndim
    bb = nodes["bbox"].astype(np.float64)
    ext = bb[:, 3:] - bb[:, :3]
    area = nodes["area"].astype(np.float64)
    cols = []
    for p in properties:
        if p in ("area", "num_pixels"):
            cols.append(area[:, None])
        elif p == "label":
            cols.append(np.ones((n, 1)))          # the callback paints the node with label=True
        elif p == "area_bbox":
            cols.append(ext.prod(axis=1)[:, None])
        elif p == "extent":
            cols.append((area / ext.prod(axis=1))[:, None])
        elif p == "bbox":
            cols.append(np.concatenate([bb[:, lo:3], bb[:, 3 + lo:]], axis=1))
        elif p == "centroid":
            cols.append(host["centroid"][:, lo:])
        else:
            cols.append(host[{"intensity_mean": "mean", "intensity_std": "std", "intensity_min": "min",
                              "intensity_max": "max"}[p]])
    return np.concatenate(cols, axis=1).astype(np.float32) if cols else np.zeros((n, 0), np.float32)
