"""`segment()` -- drop-in for ultrack.core.segmentation.processing.segment.

Same signature, config fields, SQL rows and side file as the reference
(ultrack/core/segmentation/processing.py:160-340,393-488); per timepoint the body of
`_process` (threshold, create_hierarchies, Hierarchy.nodes, ids, overlaps) is one call of
the CUDA library (`ulk_segment_frame`).  Timepoints are sharded across the ranks of the
default torch.distributed group when one exists (one process per GPU).
There is no CPU path: the call raises when the CUDA library or a GPU is missing.
"""
from __future__ import annotations

import ctypes as C
import logging
import os
import pickle
import sqlite3
import time as _time
import warnings
from typing import Dict, List, Optional

import numpy as np

from ultrack_b200 import _lib, database, frames, sharding
from ultrack_b200.config import ws_hierarchy_name
from ultrack_b200.node import dumps_node, frame_blobs

LOG = logging.getLogger(__name__)


def rename_argument(from_name: str, to_name: str):
    """ultrack/utils/deprecation.py: keeps `detection=` / `edge=` working."""
    from functools import wraps

    def _wrapper(func):
        @wraps(func)
        def _update_from_dict(*args, **kwargs):
            if from_name in kwargs:
                if to_name in kwargs:
                    raise ValueError(f"Argument {to_name} already defined, please do not mix {from_name} and {to_name} in one call.")
                warnings.warn(f"Argument {from_name} is deprecated, please use {to_name} instead.",
                              category=DeprecationWarning, stacklevel=2)
                kwargs = kwargs.copy()
                kwargs[to_name] = kwargs.pop(from_name)
            return func(*args, **kwargs)

        return _update_from_dict

    return _wrapper


def check_array_chunk(array) -> None:
    """ultrack/utils/array.py:85-96."""
    if hasattr(array, "chunks") and array.chunks is not None:
        chunk_shape = array.chunks
        if isinstance(chunk_shape[0], tuple):
            chunk_shape = chunk_shape[0]
        if chunk_shape[0] != 1:
            warnings.warn(f"Array not chunked over time dimension. Found chunk of shape {array.chunks}."
                          "Performance will be slower.")


_now = _time.perf_counter


def _generate_id(index, time: int, max_segments: int):
    """processing.py:42-59."""
    return index + (time + 1) * max_segments


def seg_params(config, fg_dtype) -> _lib.SegParams:
    """SegmentationConfig -> ulk_seg_params.  numpy evaluates `foreground > threshold` in the
    array's own float precision (processing.py:212), so the threshold is rounded to it here."""
    thr = float(config.threshold)
    dt = np.dtype(fg_dtype)
    if np.issubdtype(dt, np.floating):
        thr = float(dt.type(thr))
    name = ws_hierarchy_name(config.ws_hierarchy)
    if name not in _lib.WS_HIERARCHY_CODES:
        raise NotImplementedError(f"ws_hierarchy={name!r} is not implemented by the CUDA library")
    return _lib.SegParams(thr, float(config.min_area), float(config.max_area), float(config.min_area_factor),
                          float(config.min_frontier), _lib.WS_HIERARCHY_CODES[name], 0)


def _is_cuda_tensor(a) -> bool:
    return type(a).__module__.startswith("torch") and getattr(a, "is_cuda", False)


def segment_frame_device(lib, fg, ct, prm: _lib.SegParams, keep_labels: bool = False, stream: int = 0):
    """One `ulk_segment_frame*` call.  `fg` / `ct`: one frame, numpy (host) or torch CUDA tensors."""
    if tuple(fg.shape) != tuple(ct.shape):
        raise ValueError(f"`foreground` and `contours` shape must match. Found {fg.shape} and {ct.shape}")
    if fg.ndim not in (2, 3):
        raise ValueError(f"frames must be 2-D or 3-D, found shape {fg.shape}")
    shape3 = tuple(fg.shape) if fg.ndim == 3 else (1,) + tuple(fg.shape)
    shp = (C.c_int64 * 3)(*shape3)
    handle = C.POINTER(_lib.FrameNodes)()
    if _is_cuda_tensor(fg) and _is_cuda_tensor(ct):
        import torch

        fg, ct = fg.contiguous(), ct.contiguous()
        np_dt = lambda t: np.dtype(str(t.dtype).replace("torch.", "")) if t.dtype != torch.bool else np.dtype(bool)  # noqa: E731
        rc = lib.ulk_segment_frame(C.c_void_p(fg.data_ptr()), _lib.dtype_code(np_dt(fg)),
                                   C.c_void_p(ct.data_ptr()), _lib.dtype_code(np_dt(ct)),
                                   shp, C.byref(prm), int(keep_labels), C.c_void_p(stream), C.byref(handle))
    else:
        fg = np.ascontiguousarray(fg)
        ct = np.ascontiguousarray(ct)
        rc = lib.ulk_segment_frame_host(fg.ctypes.data_as(C.c_void_p), _lib.dtype_code(fg.dtype),
                                        ct.ctypes.data_as(C.c_void_p), _lib.dtype_code(ct.dtype),
                                        shp, C.byref(prm), int(keep_labels), C.c_void_p(stream), C.byref(handle))
    _lib.check(lib, rc)
    return handle


class FrameUpload:
    """Host->device copy of one frame in flight on the library's copy stream (`ulk_upload_begin`).

    `segment()` begins the upload of frame t+1 before it segments frame t, so the PCIe copy is
    hidden behind the kernels (the copy is asynchronous for pinned host memory only; pageable
    numpy memory still works, the copy then happens inside `begin`)."""

    def __init__(self, lib, fg, ct):
        if tuple(fg.shape) != tuple(ct.shape):
            raise ValueError(f"`foreground` and `contours` shape must match. Found {fg.shape} and {ct.shape}")
        if fg.ndim not in (2, 3):
            raise ValueError(f"frames must be 2-D or 3-D, found shape {fg.shape}")
        self.lib = lib
        self.keep = (fg, ct)          # the host buffers must outlive the copy
        shape3 = tuple(fg.shape) if fg.ndim == 3 else (1,) + tuple(fg.shape)
        self.handle = C.c_void_p()
        _lib.check(lib, lib.ulk_upload_begin(
            C.c_void_p(_host_ptr(fg)), _lib.dtype_code(_np_dtype(fg)), C.c_void_p(_host_ptr(ct)),
            _lib.dtype_code(_np_dtype(ct)), (C.c_int64 * 3)(*shape3), C.byref(self.handle)))

    def segment(self, prm: _lib.SegParams, keep_labels: bool = False, stream: int = 0):
        """`ulk_segment_frame_uploaded`: waits for the copy on `stream`, segments, releases the upload."""
        out = C.POINTER(_lib.FrameNodes)()
        handle, self.handle = self.handle, None
        rc = self.lib.ulk_segment_frame_uploaded(handle, C.byref(prm), int(keep_labels), C.c_void_p(stream), C.byref(out))
        self.keep = None
        _lib.check(self.lib, rc)
        return out

    def __del__(self):  # noqa: D105
        if getattr(self, "handle", None):
            self.lib.ulk_upload_free(self.handle)
            self.handle = None


def _np_dtype(a) -> np.dtype:
    if type(a).__module__.startswith("torch"):
        name = str(a.dtype).replace("torch.", "")
        return np.dtype(bool) if name == "bool" else np.dtype(name)
    return a.dtype


def _host_ptr(a) -> int:
    if type(a).__module__.startswith("torch"):
        if not a.is_contiguous():
            raise ValueError("host frames must be contiguous")
        return a.data_ptr()
    if not a.flags["C_CONTIGUOUS"]:
        raise ValueError("host frames must be C-contiguous")
    return a.ctypes.data


def frame_rows(nodes: Dict[str, np.ndarray], time: int, ndim: int, max_segments_per_time: int,
               features: Optional[np.ndarray] = None, lib=None, handle=None):
    """Node table -> (ids, `nodes` rows, `overlaps` rows) exactly as _process emits them
    (processing.py:235-300).  With `lib` / `handle` the pickle blobs of all nodes come from ONE
    kernel (`node.frame_blobs`) and the rows are assembled column-wise -- no per-node Python work
    beyond one tuple; without them (a table that only exists on the host) the blobs are pickled
    node by node."""
    n = len(nodes["area"])
    # the reference's running index ends at n + 1 and it raises when index > max (processing.py:305-308)
    if n + 1 > max_segments_per_time:
        raise ValueError(f"Number of segments exceeds upper bound of {max_segments_per_time} per time.")
    index = np.arange(1, n + 1, dtype=np.int64)
    ids = _generate_id(index, time, max_segments_per_time)
    parent = nodes["parent"]
    hier_parent = np.where(parent >= 0, _generate_id(parent.astype(np.int64) + 1, time, max_segments_per_time),
                           database.NO_PARENT)
    lo = 3 - ndim
    if lib is not None and handle is not None:
        buf, off = frame_blobs(lib, handle, nodes, time, ndim, (time + 1) * max_segments_per_time)
        mv = memoryview(buf)
        o = off.tolist()
        blobs = [mv[o[i]:o[i + 1]] for i in range(n)]
    else:
        blobs = []
        for i in range(n):
            bb = nodes["bbox"][i].astype(np.int64)
            blobs.append(dumps_node(int(nodes["hidx"][i]), int(ids[i]), time, int(nodes["area"][i]),
                                    float(nodes["frontier"][i]), float(nodes["height"][i]),
                                    np.concatenate([bb[lo:3], bb[3 + lo:]]), _lib.unpack_mask(nodes, i, ndim),
                                    nodes["centroid"][i][lo:].astype(np.int64)))
    cen = nodes["centroid"].astype(np.float64)
    z = cen[:, 0].tolist() if ndim == 3 else [0.0] * n
    feats = [None] * n if features is None else [pickle.dumps(f, protocol=pickle.HIGHEST_PROTOCOL) for f in features]
    rows = database.node_rows(time, ids.tolist(), hier_parent.tolist(), index.tolist(), nodes["hier_id"].tolist(), z,
                              cen[:, 1].tolist(), cen[:, 2].tolist(), nodes["area"].tolist(),
                              nodes["frontier"].tolist(), nodes["height"].tolist(), blobs, feats)
    ov = nodes["overlaps"].astype(np.int64)
    overlaps = list(zip(_generate_id(ov[:, 0] + 1, time, max_segments_per_time).tolist(),
                        _generate_id(ov[:, 1] + 1, time, max_segments_per_time).tolist()))
    return ids, rows, overlaps


def _stage(time: int, foreground, contours, config, lib):
    """Reads frame `time` (processing.py:201-212) and starts moving it to the GPU.  Returns a
    callable that segments it and the frame's ndim."""
    if config.random_seed == "frame":
        np.random.seed(time)
    fg = foreground[time]
    edge_map = contours[time]
    if config.max_noise > 0:
        # the reference draws from numpy's legacy global RNG seeded with the frame number
        # (processing.py:201-209); replayed on the host so the result stays bit-identical
        edge_map = np.asarray(edge_map.cpu() if _is_cuda_tensor(edge_map) else edge_map)
        noise = np.random.uniform(0, config.max_noise, size=edge_map.shape)
        noise = noise.astype(np.result_type(edge_map.dtype, np.float16))
        edge_map = edge_map + noise
        if _is_cuda_tensor(fg):
            fg = np.asarray(fg.cpu())
    on_device = _is_cuda_tensor(fg) and _is_cuda_tensor(edge_map)
    if not on_device:
        fg, edge_map = np.ascontiguousarray(fg), np.ascontiguousarray(edge_map)
    prm = seg_params(config, np.dtype(bool) if str(fg.dtype).endswith("bool") else str(fg.dtype).replace("torch.", ""))
    if on_device:
        return (lambda: segment_frame_device(lib, fg, edge_map, prm)), fg.ndim
    up = FrameUpload(lib, fg, edge_map)
    return (lambda: up.segment(prm)), fg.ndim


# wall-clock split of _process, accumulated per process (bench.py's `e2e_api` leg reads it)
TIMERS: Dict[str, float] = {"kernels": 0.0, "d2h": 0.0, "rows": 0.0, "sql": 0.0}


def _process(time: int, staged, db_path: str, max_segments_per_time: int,
             writer: "_RowWriter", lib, images=None, properties: Optional[List[str]] = None) -> None:
    """processing.py:160-340 for one timepoint (`staged` comes from `_stage`)."""
    run, ndim = staged
    t0 = _now()
    handle = run()
    t1 = _now()
    TIMERS["kernels"] += t1 - t0
    LOG.info(f"Computing nodes of time {time}")
    try:
        nodes = _lib.nodes_to_host(lib, handle, with_masks=False)
        t2 = _now()
        TIMERS["d2h"] += t2 - t1
        feats = None
        if properties is not None:
            from ultrack_b200 import features
            frame = None if images is None else images[time]
            if frame is not None and not _is_cuda_tensor(frame):
                frame = np.asarray(frame)
            feats = features.frame_features(lib, handle, nodes, ndim, frame, properties)
        ids, rows, overlaps = frame_rows(nodes, time, ndim, max_segments_per_time, feats, lib, handle)
        t3 = _now()
        TIMERS["rows"] += t3 - t2
    except Exception:
        lib.ulk_frame_nodes_free(handle)
        raise
    frames.STORE.put(db_path, time, frames.DeviceFrame(lib, handle, ids, ndim))
    writer.put(time, rows, overlaps)


class _RowWriter:
    """Inserts the rows of finished frames from a background thread, so that the sqlite insert of
    frame t (the largest host cost of `segment()`: tens of MB of blobs per frame) overlaps the
    upload and the kernels of frame t + 1.  sqlite3 and ctypes both release the GIL while they
    work.  One frame is queued at most (its blob buffer stays alive in the queue); errors surface
    in `close()` / the next `put()` with the reference's messages (processing.py:320-334)."""

    def __init__(self, db_path: str, catch_duplicates: bool):
        import queue
        import threading
        self.db_path, self.catch = db_path, catch_duplicates
        self.q: "queue.Queue" = queue.Queue(maxsize=1)
        self.error: Optional[BaseException] = None
        self.thread = threading.Thread(target=self._run, name="ultrack_b200-row-writer", daemon=True)
        self.thread.start()

    def _run(self) -> None:
        while True:
            item = self.q.get()
            if item is None:
                return
            time, rows, overlaps = item
            if self.error is not None:
                continue                      # drain: a previous frame failed
            t0 = _now()
            try:
                database.insert_nodes(self.db_path, rows, overlaps)
                LOG.info(f"DONE with time {time}.")
            except sqlite3.IntegrityError as e:
                frames.STORE.drop(self.db_path, time)
                if self.catch:
                    LOG.warning(f"IntegrityError: {e}\nSkipping batch {time}.\n"
                                "Cannot guarantee image was entirely processed.\nSEGMENTS MIGHT BE MISSING!")
                else:
                    err = ValueError("Duplicated nodes found. Set `overwrite=True` to overwrite existing data.")
                    err.__cause__ = e
                    self.error = err
            except BaseException as e:  # noqa: BLE001
                frames.STORE.drop(self.db_path, time)
                self.error = e
            TIMERS["sql"] += _now() - t0

    def put(self, time: int, rows, overlaps) -> None:
        if self.error is not None:
            self.close()
        self.q.put((time, rows, overlaps))

    def close(self) -> None:
        if self.thread.is_alive():
            self.q.put(None)
            self.thread.join()
        if self.error is not None:
            err, self.error = self.error, None
            raise err


@rename_argument("detection", "foreground")
@rename_argument("edge", "contours")
def segment(
    foreground,
    contours,
    config,
    max_segments_per_time: int = 1_000_000,
    batch_index: Optional[int] = None,
    overwrite: bool = False,
    insertion_throttle_rate: int = 50,
    images=None,
    properties: Optional[List[str]] = None,
) -> None:
    """Add candidate segmentation (nodes) from `foreground` and `contours` to database.

    Parameters are those of ultrack.segment (processing.py:393-430).  `insertion_throttle_rate`
    is accepted and unused: rows are inserted once per frame.  `properties`: the regionprops
    features that are reductions over the node masks (`ultrack_b200.features.SUPPORTED`) are
    computed on the device for all nodes of a frame at once; others raise NotImplementedError.
    """
    LOG.info(f"Adding nodes with SegmentationConfig:\n{config.segmentation_config}")
    if tuple(foreground.shape) != tuple(contours.shape):
        raise ValueError(f"`foreground` and `contours` shape must match. Found {foreground.shape} and {contours.shape}")
    from ultrack_b200 import features
    prop_names = features.properties_names(foreground.shape, images, properties)
    check_array_chunk(foreground)
    check_array_chunk(contours)

    lib = _lib.get_lib()
    rank, world = sharding.dist_info()
    if world > 1:
        _lib.check(lib, lib.ulk_set_device(sharding.local_rank()))
    length = foreground.shape[0]
    time_points = sharding.batch_index_range(length, config.segmentation_config.n_workers, batch_index)
    LOG.info(f"Segmenting time points {time_points}")
    db_path = config.data_config.database_path
    batch_start_index = int(os.getenv("ULTRACK_BATCH_INDEX_START", "0"))
    if rank == 0 and (batch_index is None or batch_index == batch_start_index):
        if overwrite:
            database.clear_all_data(db_path)
        database.create_all(db_path)
        meta = {"shape": [int(s) for s in foreground.shape]}
        if prop_names is not None:       # toml drops None values (processing.py:460-467)
            meta["properties"] = prop_names
        config.data_config.metadata_add(meta)
    if overwrite:
        frames.STORE.clear(db_path)
    sharding.barrier()
    # the host->device copy of the next frame overlaps the kernels of the current one
    mine = sharding.owned(time_points, length, world, rank)
    staged = _stage(mine[0], foreground, contours, config.segmentation_config, lib) if mine else None
    writer = _RowWriter(db_path, catch_duplicates=batch_index is not None)
    try:
        for k, time in enumerate(mine):
            nxt = _stage(mine[k + 1], foreground, contours, config.segmentation_config, lib) if k + 1 < len(mine) else None
            _process(time, staged, db_path, max_segments_per_time, writer, lib, images, properties)
            staged = nxt
    finally:
        writer.close()
    sharding.barrier()
