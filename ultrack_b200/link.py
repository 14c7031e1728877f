"""`link()` -- drop-in for ultrack.core.linking.processing.link.

Same signature, config fields and `links` rows as the reference
(ultrack/core/linking/processing.py:107-338); per frame pair the body of
`compute_spatial_neighbors` (KD-tree k-NN, mask IoU, top-k) is one call of the CUDA
library (`ulk_link_frames`) on device-resident node tables.  Frame pairs are sharded by
source timepoint across the ranks of the default torch.distributed group; the one table a
rank does not own -- the first frame of the next rank's block -- is received from that
rank over NCCL (frames.send_frame / recv_frame).
"""
from __future__ import annotations

import ctypes as C
import logging
import warnings
from typing import Optional, Sequence, Tuple

import numpy as np

from ultrack_b200 import _lib, database, frames, sharding

LOG = logging.getLogger(__name__)


def link_params(config, scale: Optional[Sequence[float]], ndim: int, weight_func: str = "iou",
                edge_must_be_positive: bool = False) -> _lib.LinkParams:
    """LinkingConfig (+ scale) -> ulk_link_params; `scale[-ndim:]` as in processing.py:215-219.
    `weight_func` / `edge_must_be_positive`: the two extra arguments of compute_spatial_neighbors
    (processing.py:183-184), "iou" = Node.IoU, "intersection" = Node.intersection."""
    sc = [1.0, 1.0, 1.0]
    if scale is not None:
        # the reference keeps the last min(n_dim, len(scale)) axes of BOTH position arrays
        # (processing.py:215-219): a short scale drops the leading axes from the distance, which
        # a zero scale factor on those axes reproduces exactly
        m = min(ndim, len(scale))
        sc = [0.0] * (3 - m) + [float(v) for v in list(scale)[-m:]] if m > 0 else [0.0, 0.0, 0.0]
    prm = _lib.LinkParams(float(config.max_distance), float(config.distance_weight), int(config.max_neighbors),
                          int(scale is not None), (C.c_double * 3)(*sc))
    prm.weight_kind = _lib.WEIGHT_KIND_CODES[weight_func]
    prm.edge_must_be_positive = int(bool(edge_must_be_positive))
    return prm


def frame_color_stats(lib, frame: frames.DeviceFrame, images: Sequence, time: int):
    """Per-node intensity mean / std of every image channel at `time` (_compute_features,
    processing.py:31-49) on the device node table; cached on the frame (a frame is the source of
    one pair and the target of the previous one, with the same image)."""
    key = (tuple(id(im) for im in images), int(time))
    cached = getattr(frame, "_color", None)
    if cached is not None and cached[0] == key:
        return cached[1]
    from ultrack_b200 import capi
    chans = []
    for im in images:
        a = im[time]
        a = a if type(a).__module__.startswith("torch") else np.asarray(a)
        if a.ndim == frame.ndim + 1 and a.shape[-1] == 1:
            a = a[..., 0]
        if a.ndim != frame.ndim:
            raise ValueError(f"image frame of shape {tuple(a.shape)} does not match {frame.ndim}-D nodes; "
                             "pass one array per channel")
        chans.append(a)
    st = capi.node_stats(lib, frame.handle, chans)
    frame._color = (key, st)
    return st


def link_frames_device(lib, src: frames.DeviceFrame, dst: frames.DeviceFrame, prm: _lib.LinkParams,
                       shift: Optional[np.ndarray] = None, stream: int = 0,
                       color=None, z_score_threshold: float = 5.0) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """One `ulk_link_frames` call -> (weight, source_id, target_id) rows.  `color` = (source stats,
    target stats) from `frame_color_stats` switches the colour z-score gate on."""
    if color is not None:
        prm.n_channels = color[0]["mean"].shape[1]
        prm.z_score_threshold = float(z_score_threshold)
        prm.src_mean, prm.dst_mean, prm.dst_std = color[0]["mean"].ptr, color[1]["mean"].ptr, color[1]["std"].ptr
    out = C.POINTER(_lib.Links)()
    sh = None
    if shift is not None:
        shift = np.array(shift, dtype=np.float64, order="C")
        if dst.ndim == 2:
            shift[:, 0] = 0.0      # `target_shift[:, -n_dim:]` (processing.py:212): no z axis in 2-D
        sh = shift.ctypes.data_as(C.c_void_p)
    _lib.check(lib, lib.ulk_link_frames(src.handle, dst.handle, sh, C.byref(prm), C.c_void_p(stream), C.byref(out)))
    try:
        n = int(out.contents.n_links)
        s = np.zeros(n, np.int32)
        t = np.zeros(n, np.int32)
        w = np.zeros(n, np.float64)
        p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        _lib.check(lib, lib.ulk_links_to_host(out, p(s), p(t), p(w), C.c_void_p(stream)))
    finally:
        lib.ulk_links_free(out)
    return w, src.ids[s], dst.ids[t]


def _frame(lib, db: str, t: int, shape3, ndim: int) -> Optional[frames.DeviceFrame]:
    f = frames.STORE.get(db, t)
    if f is None:
        if sharding.dist_info()[1] > 1:
            LOG.warning(f"node table of t={t} is not resident on rank {sharding.dist_info()[0]}: rebuilding it from "
                        "the database blobs (slow path; segment() and link() should run in the same processes)")
        f, _ = frames.frame_from_database(lib, db, t, shape3, ndim)
        if f is not None:
            frames.STORE.put(db, t, f)
    return f


def _process(time: int, config, db_path: str, scale, lib, shape3, ndim: int,
             next_frame: Optional[frames.DeviceFrame] = None, images: Sequence = ()) -> None:
    """linking/processing.py:107-169 + 172-291 for the pair (time, time + 1)."""
    src = _frame(lib, db_path, time, shape3, ndim)
    dst = next_frame if next_frame is not None else _frame(lib, db_path, time + 1, shape3, ndim)
    if dst is None or dst.n_nodes == 0:
        print(f"No nodes found for time {time + 1}")
        warnings.warn(f"No segments found at target t={time + 1}; skipping linking from t={time}.",
                      RuntimeWarning, stacklevel=2)
        return
    if src is None or src.n_nodes == 0:
        print(f"No nodes found for time {time}")
        warnings.warn(f"No segments found at source t={time}; skipping linking to t={time + 1}.",
                      RuntimeWarning, stacklevel=2)
        return
    shift = frames.load_shifts(db_path, time + 1, dst.n_nodes)
    prm = link_params(config, scale, ndim)
    color = None
    if len(images) > 0:
        LOG.info(f"computing filtering by color z-score from t={time}")
        color = (frame_color_stats(lib, src, images, time), frame_color_stats(lib, dst, images, time + 1))
    weight, source_id, target_id = link_frames_device(lib, src, dst, prm, shift, color=color,
                                                      z_score_threshold=config.z_score_threshold)
    if len(weight) == 0:
        LOG.warning(f"No links found for time {time}. Increase `linking_config.max_distance` parameter.")
        return
    LOG.info(f"Pushing links from time {time} to {db_path}")
    database.insert_links(db_path, weight, source_id, target_id)


def link(
    config,
    images: Sequence = tuple(),
    scale: Optional[Sequence[float]] = None,
    batch_index: Optional[int] = None,
    overwrite: bool = False,
) -> None:
    """Links candidate segments (nodes) with their neighbors on the next time.

    Parameters are those of ultrack.link (linking/processing.py:294-318).  `images`: one array
    (T, (Z), Y, X) per channel; per-node intensity mean / std are reduced over the device masks
    (`ulk_node_stats`) and candidates outside `z_score_threshold` are dropped (processing.py:52-104).
    """
    LOG.info(f"Linking nodes with LinkingConfig:\n{config.linking_config}")
    lib = _lib.get_lib()
    rank, world = sharding.dist_info()
    if world > 1:
        _lib.check(lib, lib.ulk_set_device(sharding.local_rank()))
    db_path = config.data_config.database_path
    max_t = database.maximum_time_from_database(db_path)
    time_points = sharding.batch_index_range(max_t, config.linking_config.n_workers, batch_index)
    LOG.info(f"Linking time points {time_points}")
    if overwrite and (batch_index is None or batch_index == 0) and rank == 0:
        database.clear_linking_data(db_path)
    sharding.barrier()
    shape = config.data_config.metadata.get("shape")
    if shape is None:
        raise ValueError(f"`shape` missing from {config.data_config.metadata_path}; run segment() first")
    ndim = len(shape) - 1
    shape3 = tuple(shape[1:]) if ndim == 3 else (1,) + tuple(shape[1:])

    # pairs follow their SOURCE frame under segment()'s partition of the max_t + 1 frames
    mine = sharding.owned(time_points, max_t + 1, world, rank)
    boundary = None
    if world > 1:
        boundary = _exchange_boundary(lib, db_path, time_points, max_t + 1, world, rank, shape3, ndim)
    for time in mine:
        nxt = boundary if (boundary is not None and time == mine[-1]) else None
        _process(time, config.linking_config, db_path, scale, lib, shape3, ndim, next_frame=nxt, images=images)
    sharding.barrier()


def _exchange_boundary(lib, db_path: str, time_points: Sequence[int], total: int, world: int, rank: int, shape3, ndim: int):
    """The target frame of a rank's last pair belongs to the next rank's block: its owner sends
    the node table (frames.send_frame / recv_frame, NCCL over NVLink on a GPU group).  All ranks
    walk the same schedule, so sends and receives pair up without a handshake."""
    received = None
    for r in range(world):
        pairs = sharding.owned(time_points, total, world, r)
        if not pairs:
            continue
        target = pairs[-1] + 1
        sender = sharding.owner_of(target, total, world)
        if sender == r:
            continue
        if rank == sender:
            frames.send_frame(_frame(lib, db_path, target, shape3, ndim), r)
        elif rank == r:
            received = frames.recv_frame(lib, sender)
    return received
