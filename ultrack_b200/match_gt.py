"""Ground-truth candidate links: the `compute_spatial_neighbors` call of the reference's GT matcher.

`ultrack.core.match_gt._match_ground_truth_frame` (core/match_gt.py:43-185) turns the objects
of one ground-truth label frame into nodes, links every GT node (as *target*) to the candidate
segments of the same timepoint (as *sources*) with `weight_func = Node.intersection` and
`edge_must_be_positive=True` (match_gt.py:148-168), writes those rows to `gt_links` and then
solves an assignment ILP (`SQLGTMatcher`).  The k-NN + mask-intersection part is the same
kernel as linking (`ulk_link_frames` with `weight_kind = intersection`); this module runs it
on the device-resident node table of the frame and fills `gt_nodes` / `gt_links`.  The ILP that
follows is the solver side of ultrack and out of scope.
"""
from __future__ import annotations

import logging
from typing import Optional, Sequence, Tuple

import numpy as np
import scipy.ndimage as ndi

from ultrack_b200 import _lib, database, frames
from ultrack_b200.link import _frame, link_frames_device, link_params
from ultrack_b200.node import dumps_node

LOG = logging.getLogger(__name__)


def gt_frame_nodes(gt_labels: np.ndarray):
    """`regionprops(gt_labels)` + `Node.from_mask` (match_gt.py:79-101): labels ascending, bbox
    (mins, exclusive maxes), bool mask over the bbox, centroid = round-half-even of the float64
    mean coordinate (core/segmentation/node.py:229-233)."""
    gt_labels = np.asarray(gt_labels)
    ndim = gt_labels.ndim
    labels, bboxes, cens, masks, areas = [], [], [], [], []
    for lab, sl in enumerate(ndi.find_objects(gt_labels.astype(np.int64, copy=False)), start=1):
        if sl is None:
            continue
        m = gt_labels[sl] == lab
        lo = np.array([s.start for s in sl], np.int64)
        hi = np.array([s.stop for s in sl], np.int64)
        c = np.round(np.mean(np.nonzero(m), axis=1) + lo).astype(np.int64)
        labels.append(lab)
        bboxes.append(np.concatenate([lo, hi]))
        cens.append(c)
        masks.append(m)
        areas.append(int(m.sum()))
    return ndim, np.array(labels, np.int64), bboxes, cens, masks, np.array(areas, np.int64)


def match_ground_truth_links(time: int, gt_labels, config, scale: Optional[Sequence[float]] = None,
                             write: bool = True) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """GT nodes of `gt_labels[time]` linked to the candidate segments of `time`:
    returns (weight, source_id = candidate node id, target_id = GT label) and, with `write`,
    appends the `gt_nodes` / `gt_links` rows (match_gt.py:103-168)."""
    lib = _lib.get_lib()
    db_path = config.data_config.database_path
    ndim, labels, bboxes, cens, masks, areas = gt_frame_nodes(np.asarray(gt_labels[time]))
    empty = (np.zeros(0), np.zeros(0, np.int64), np.zeros(0, np.int64))
    if len(labels) == 0:
        LOG.warning(f"No objects found in time point {time}")
        return empty
    shape = config.data_config.metadata.get("shape")
    shape3 = tuple(shape[1:]) if ndim == 3 else (1,) + tuple(shape[1:])
    src = _frame(lib, db_path, time, shape3, ndim)
    if src is None or src.n_nodes == 0:
        return empty
    n = len(labels)
    bbox3 = np.zeros((n, 6), np.int32)
    cen3 = np.zeros((n, 3), np.int32)
    for i in range(n):
        bbox3[i, 3 - ndim:3] = bboxes[i][:ndim]
        bbox3[i, 6 - ndim:] = bboxes[i][ndim:]
        bbox3[i, 3:6 - ndim] = 1
        cen3[i, 3 - ndim:] = cens[i]
    off, words = frames.pack_masks(masks, bbox3)
    gt = frames.upload(lib, shape3, bbox3, cen3, areas.astype(np.int32), off, words, labels, ndim)
    try:
        prm = link_params(config.linking_config, scale, ndim, weight_func="intersection", edge_must_be_positive=True)
        weight, source_id, target_id = link_frames_device(lib, src, gt, prm)
    finally:
        gt.free()
    if write:
        rows = []
        for i in range(n):
            blob = dumps_node(-1, int(labels[i]), time, int(areas[i]), -1.0, -1.0, bboxes[i], masks[i], cens[i])
            z = int(cens[i][0]) if ndim == 3 else 0
            rows.append((int(time), int(labels[i]), blob, float(z), float(cens[i][-2]), float(cens[i][-1])))
        database.insert_gt(db_path, rows, weight, source_id, target_id)
    return weight, source_id, target_id
