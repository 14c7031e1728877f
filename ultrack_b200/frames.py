"""Device-resident node tables of the frames of a database.

`segment()` leaves every frame's node table (bbox, centroid, area, bit-packed masks) on
the GPU so that `link()` never re-reads the `nodes.pickle` blobs the way the reference
does (ultrack/core/linking/processing.py:136-148 unpickles every mask twice per frame).
When a table is not resident -- `link()` called from a fresh process, or on a database
written by the reference -- it is rebuilt from the blobs and uploaded once.
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import numpy as np

from ultrack_b200 import _lib, database
from ultrack_b200.node import loads_node


def _vp(a: Optional[np.ndarray]):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class DeviceFrame:
    """Owns one `ulk_frame_nodes` handle plus the SQL ids of its rows."""

    def __init__(self, lib, handle, ids: np.ndarray, ndim: int):
        self.lib = lib
        self.handle = handle
        self.ids = ids
        self.ndim = ndim

    @property
    def n_nodes(self) -> int:
        return int(self.handle.contents.n_nodes)

    @property
    def nbytes(self) -> int:
        f = self.handle.contents
        return int(f.n_nodes) * 64 + int(f.mask_words) * 4

    def free(self) -> None:
        if self.handle is not None:
            self.lib.ulk_frame_nodes_free(self.handle)
            self.handle = None

    def __del__(self):  # noqa: D105
        try:
            self.free()
        except Exception:  # noqa: BLE001
            pass


class FrameStore:
    """(database path, t) -> DeviceFrame, least-recently-used eviction above `max_bytes`."""

    def __init__(self, max_bytes: int = 64 << 30):
        self.max_bytes = max_bytes
        self._sized = False
        self._frames: "OrderedDict[Tuple[str, int], DeviceFrame]" = OrderedDict()
        self._bytes = 0

    def _size_from_device(self) -> None:
        """Caps the store at half of what is free on the device once the first table exists (the
        scratch slab of the segmentation call is reserved by then)."""
        self._sized = True
        try:
            import torch
            if torch.cuda.is_available():
                free, _ = torch.cuda.mem_get_info()
                self.max_bytes = min(self.max_bytes, int(free // 2))
        except Exception:  # noqa: BLE001
            pass

    def put(self, db: str, t: int, frame: DeviceFrame) -> None:
        if not self._sized:
            self._size_from_device()
        self.drop(db, t)
        self._frames[(db, t)] = frame
        self._bytes += frame.nbytes
        while self._bytes > self.max_bytes and len(self._frames) > 2:
            _, old = self._frames.popitem(last=False)
            self._bytes -= old.nbytes
            old.free()

    def get(self, db: str, t: int) -> Optional[DeviceFrame]:
        f = self._frames.get((db, t))
        if f is not None:
            self._frames.move_to_end((db, t))
        return f

    def drop(self, db: str, t: int) -> None:
        f = self._frames.pop((db, t), None)
        if f is not None:
            self._bytes -= f.nbytes
            f.free()

    def clear(self, db: Optional[str] = None) -> None:
        for key in [k for k in self._frames if db is None or k[0] == db]:
            self.drop(*key)


STORE = FrameStore()


def pack_masks(masks, bboxes3: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Bool masks -> the library's layout: rows packed along x, 32 voxels per little-endian word."""
    n = len(masks)
    off = np.zeros(n + 1, np.int64)
    ext = (bboxes3[:, 3:] - bboxes3[:, :3]).astype(np.int64)
    off[1:] = np.cumsum(ext[:, 0] * ext[:, 1] * ((ext[:, 2] + 31) // 32))
    words = np.zeros(int(off[-1]), np.uint32)
    for i, m in enumerate(masks):
        m3 = np.asarray(m, dtype=bool).reshape(tuple(ext[i]))
        wpr = (m3.shape[2] + 31) // 32
        padded = np.zeros(m3.shape[:2] + (wpr * 32,), np.uint8)
        padded[:, :, :m3.shape[2]] = m3
        words[off[i]:off[i + 1]] = np.packbits(padded, axis=-1, bitorder="little").view(np.uint32).ravel()
    return off, words


def upload(lib, shape3, bbox3: np.ndarray, centroid3: np.ndarray, area: np.ndarray, mask_off: np.ndarray,
           mask: np.ndarray, ids: np.ndarray, ndim: int, stream: int = 0) -> DeviceFrame:
    handle = C.POINTER(_lib.FrameNodes)()
    shp = (C.c_int32 * 3)(*[int(s) for s in shape3])
    bbox3 = np.ascontiguousarray(bbox3, np.int32)
    centroid3 = np.ascontiguousarray(centroid3, np.int32)
    area = np.ascontiguousarray(area, np.int32)
    mask_off = np.ascontiguousarray(mask_off, np.int64)
    mask = np.ascontiguousarray(mask, np.uint32)
    _lib.check(lib, lib.ulk_frame_nodes_from_host(len(ids), shp, _vp(bbox3), _vp(centroid3), _vp(area),
                                                  _vp(mask_off), _vp(mask), C.c_void_p(stream), C.byref(handle)))
    return DeviceFrame(lib, handle, np.asarray(ids, np.int64), ndim)


def frame_from_database(lib, db: str, t: int, shape3, ndim: int) -> Tuple[Optional[DeviceFrame], Optional[np.ndarray]]:
    """Rebuilds the node table of frame `t` from the `nodes` rows (ids ascending) and
    returns it with the (z,y,x) shift columns."""
    rows = database.load_frame_nodes(db, t)
    if len(rows) == 0:
        return None, None
    n = len(rows)
    ids = np.array([r[0] for r in rows], np.int64)
    shifts = np.array([[r[2] or 0.0, r[3] or 0.0, r[4] or 0.0] for r in rows], np.float64)
    bbox3 = np.zeros((n, 6), np.int32)
    cen3 = np.zeros((n, 3), np.int32)
    area = np.zeros(n, np.int32)
    masks = []
    for i, r in enumerate(rows):
        node = loads_node(r[1])
        bb = np.asarray(node.bbox, np.int64)
        d = len(bb) // 2
        bbox3[i, 3 - d:3] = bb[:d]
        bbox3[i, 6 - d:] = bb[d:]
        bbox3[i, 3:6 - d] = 1
        cen3[i, 3 - d:] = np.asarray(node.centroid, np.int64)
        m = np.asarray(node.mask, bool)
        # the reference recounts mask.sum() for the IoU (node.py:116); `area` equals it
        area[i] = int(m.sum())
        masks.append(m)
    off, words = pack_masks(masks, bbox3)
    return upload(lib, shape3, bbox3, cen3, area, off, words, ids, ndim), shifts


def load_shifts(db: str, t: int, n: int) -> Optional[np.ndarray]:
    """(z,y,x)_shift columns of frame `t` ordered by id, or None when all are zero."""
    with database.connect(db) as conn:
        rows = conn.execute("SELECT z_shift, y_shift, x_shift FROM nodes WHERE t = ? ORDER BY id", (int(t),)).fetchall()
    sh = np.array([[a or 0.0, b or 0.0, c or 0.0] for a, b, c in rows], np.float64).reshape(-1, 3)
    if sh.shape[0] != n:
        raise ValueError(f"frame {t}: {sh.shape[0]} rows in the database, {n} nodes on the device")
    return sh if np.any(sh != 0.0) else None


# ---------------------------------------------------------------------- shard-boundary exchange
def _dist_device():
    import torch
    import torch.distributed as dist

    if dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def send_frame(frame: Optional[DeviceFrame], dst: int) -> None:
    """Sends a node table to rank `dst` (NCCL over NVLink when the group is nccl): header,
    ids + mask offsets (int64), bbox | centroid | area | mask words (int32)."""
    import torch
    import torch.distributed as dist

    dev = _dist_device()
    n = frame.n_nodes if frame is not None else 0
    words = int(frame.handle.contents.mask_words) if n else 0
    shape = list(frame.handle.contents.shape) if frame is not None else [0, 0, 0]
    ndim = frame.ndim if frame is not None else 0
    header = torch.tensor([n, words, ndim] + [int(s) for s in shape], dtype=torch.int64, device=dev)
    dist.send(header, dst)
    if n == 0:
        return
    i64buf = torch.empty(2 * n + 1, dtype=torch.int64, device=dev)
    i32buf = torch.empty(10 * n + words, dtype=torch.int32, device=dev)
    i64buf[:n] = torch.from_numpy(frame.ids).to(dev)
    base32, base64 = i32buf.data_ptr(), i64buf.data_ptr()
    _lib.check(frame.lib, frame.lib.ulk_frame_nodes_to_device(
        frame.handle, C.c_void_p(base32), C.c_void_p(base32 + 24 * n), C.c_void_p(base32 + 36 * n),
        C.c_void_p(base64 + 8 * n), C.c_void_p(base32 + 40 * n) if words else None, None))
    dist.send(i64buf, dst)
    dist.send(i32buf, dst)


def recv_frame(lib, src: int) -> Optional[DeviceFrame]:
    import torch
    import torch.distributed as dist

    dev = _dist_device()
    header = torch.empty(6, dtype=torch.int64, device=dev)
    dist.recv(header, src)
    n, words, ndim, d, h, w = [int(v) for v in header.cpu().tolist()]
    if n == 0:
        return None
    i64buf = torch.empty(2 * n + 1, dtype=torch.int64, device=dev)
    i32buf = torch.empty(10 * n + words, dtype=torch.int32, device=dev)
    dist.recv(i64buf, src)
    dist.recv(i32buf, src)
    if dev.type == "cuda":
        torch.cuda.synchronize()
    base32, base64 = i32buf.data_ptr(), i64buf.data_ptr()
    handle = C.POINTER(_lib.FrameNodes)()
    shp = (C.c_int32 * 3)(d, h, w)
    _lib.check(lib, lib.ulk_frame_nodes_from_device(
        n, shp, C.c_void_p(base32), C.c_void_p(base32 + 24 * n), C.c_void_p(base32 + 36 * n),
        C.c_void_p(base64 + 8 * n), words, C.c_void_p(base32 + 40 * n) if words else None, None, C.byref(handle)))
    return DeviceFrame(lib, handle, i64buf[:n].cpu().numpy().copy(), ndim)
