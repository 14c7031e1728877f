"""Thin Python calls of the C ABI on host (numpy) arrays: what bench.py, smoke() and the tests
share.  `lib` is the ctypes handle from `_lib.get_lib()` (the tests may pass the host-emulation
harness instead, whose `ulk_segment_frame` takes host pointers)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np

from ultrack_b200 import _lib


def seg_params(threshold=0.5, min_area=100, max_area=1_000_000, min_area_factor=4.0,
               min_frontier=0.0, ws_hierarchy="area", fg_dtype=None) -> _lib.SegParams:
    thr = float(threshold)
    if fg_dtype is not None and np.issubdtype(np.dtype(fg_dtype), np.floating):
        thr = float(np.dtype(fg_dtype).type(threshold))
    return _lib.SegParams(thr, float(min_area), float(max_area), float(min_area_factor),
                          float(min_frontier), _lib.WS_HIERARCHY_CODES[ws_hierarchy], 0)


def segment_frame(lib, fg: np.ndarray, ct: np.ndarray, keep_labels: bool = True, **kw):
    """Segments one frame held in host (numpy) arrays; returns (handle, numpy node table).
    The emulation library takes host pointers in `ulk_segment_frame`; the CUDA library takes
    them in `ulk_segment_frame_host` (copies are part of the call)."""
    fg = np.ascontiguousarray(fg)
    ct = np.ascontiguousarray(ct)
    shape = fg.shape if fg.ndim == 3 else (1,) + fg.shape
    shp = (C.c_int64 * 3)(*shape)
    prm = seg_params(fg_dtype=fg.dtype, **kw)
    handle = C.POINTER(_lib.FrameNodes)()
    entry = lib.ulk_segment_frame if getattr(lib, "_ulk_is_emu", False) else lib.ulk_segment_frame_host
    rc = entry(fg.ctypes.data_as(C.c_void_p), _lib.dtype_code(fg.dtype),
               ct.ctypes.data_as(C.c_void_p), _lib.dtype_code(ct.dtype),
               shp, C.byref(prm), int(keep_labels), None, C.byref(handle))
    _lib.check(lib, rc)
    return handle, _lib.nodes_to_host(lib, handle, with_labels=keep_labels)


def link_frames(lib, src, dst, max_distance=15.0, max_neighbors=5, distance_weight=0.0,
                shift: Optional[np.ndarray] = None, scale: Optional[Sequence[float]] = None):
    prm = _lib.LinkParams(float(max_distance), float(distance_weight), int(max_neighbors),
                          int(scale is not None), (C.c_double * 3)(*(scale if scale is not None else (1, 1, 1))))
    out = C.POINTER(_lib.Links)()
    sh = None
    if shift is not None:
        shift = np.ascontiguousarray(shift, dtype=np.float64)
        sh = shift.ctypes.data_as(C.c_void_p)
    _lib.check(lib, lib.ulk_link_frames(src, dst, sh, C.byref(prm), None, C.byref(out)))
    n = int(out.contents.n_links)
    s = np.zeros(n, np.int32)
    t = np.zeros(n, np.int32)
    w = np.zeros(n, np.float64)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    _lib.check(lib, lib.ulk_links_to_host(out, p(s), p(t), p(w), None))
    npairs = int(out.contents.n_pairs)
    lib.ulk_links_free(out)
    return s, t, w, npairs
