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


class DeviceArray:
    """A float64 [n, c] array the library can read: a torch CUDA tensor for the CUDA library, a
    numpy array for the host-emulation harness (whose "device" pointers are host pointers)."""

    def __init__(self, lib, n: int, c: int):
        self.shape = (int(n), int(c))
        if getattr(lib, "_ulk_is_emu", False):
            self._np = np.zeros(self.shape, np.float64)
            self._t = None
            self.ptr = self._np.ctypes.data
        else:
            import torch
            self._np = None
            self._t = torch.zeros(self.shape, dtype=torch.float64, device="cuda")
            self.ptr = self._t.data_ptr()

    def host(self) -> np.ndarray:
        return self._np if self._np is not None else self._t.cpu().numpy()


def _image_to_lib(lib, image):
    """(keepalive, pointer) of one [D,H,W] image channel where the library can read it."""
    if type(image).__module__.startswith("torch") and image.is_cuda:
        t = image.contiguous()
        return t, t.data_ptr()
    image = np.ascontiguousarray(image)
    if getattr(lib, "_ulk_is_emu", False):
        return image, image.ctypes.data
    import torch
    src = image.view(np.int16) if image.dtype == np.uint16 else (image.view(np.int32) if image.dtype == np.uint32 else image)
    t = torch.from_numpy(src).cuda()
    return t, t.data_ptr()


def node_stats(lib, handle, images: Sequence[np.ndarray] = (), stream: int = 0, centroid: bool = False,
               minmax: bool = False):
    """`ulk_node_stats` over every image channel -> dict of DeviceArrays: "mean", "std" [n, c]
    (core/segmentation/node.py:280-296), optionally "min", "max" [n, c] and "centroid" [n, 3]."""
    n, c = int(handle.contents.n_nodes), len(images)
    out = {}
    if c:
        out["mean"], out["std"] = DeviceArray(lib, n, c), DeviceArray(lib, n, c)
        if minmax:
            out["min"], out["max"] = DeviceArray(lib, n, c), DeviceArray(lib, n, c)
    if centroid:
        out["centroid"] = DeviceArray(lib, n, 3)
    ptr = lambda k: C.c_void_p(out[k].ptr) if k in out else None  # noqa: E731
    for ch in range(max(c, 1)):
        keep, img, code = None, None, 0
        if c:
            a = images[ch]
            if type(a).__module__.startswith("torch"):
                dt = np.dtype(bool) if str(a.dtype) == "torch.bool" else np.dtype(str(a.dtype).replace("torch.", ""))
                a = a if a.is_cuda else a.numpy()
            else:
                a = np.asarray(a)
                dt = a.dtype
            keep, p = _image_to_lib(lib, a)
            img, code = C.c_void_p(p), _lib.dtype_code(dt)
        _lib.check(lib, lib.ulk_node_stats(handle, img, code, ch, max(c, 1), ptr("mean"), ptr("std"), ptr("min"),
                                           ptr("max"), ptr("centroid") if ch == 0 else None, C.c_void_p(stream)))
        del keep
    return out


def node_intensity(lib, handle, images: Sequence[np.ndarray], stream: int = 0):
    """(mean, std) of `node_stats`: the inputs of the colour gate of `ulk_link_frames`."""
    st = node_stats(lib, handle, images, stream)
    return st["mean"], st["std"]


def link_frames(lib, src, dst, max_distance=15.0, max_neighbors=5, distance_weight=0.0,
                shift: Optional[np.ndarray] = None, scale: Optional[Sequence[float]] = None,
                weight_func: str = "iou", edge_must_be_positive: bool = False,
                color=None, z_score_threshold: float = 5.0):
    """`color` = (src_mean, dst_mean, dst_std) DeviceArrays from `node_intensity` (colour gate)."""
    prm = _lib.LinkParams(float(max_distance), float(distance_weight), int(max_neighbors),
                          int(scale is not None), (C.c_double * 3)(*(scale if scale is not None else (1, 1, 1))))
    prm.weight_kind = _lib.WEIGHT_KIND_CODES[weight_func]
    prm.edge_must_be_positive = int(bool(edge_must_be_positive))
    if color is not None:
        prm.n_channels = color[0].shape[1]
        prm.z_score_threshold = float(z_score_threshold)
        prm.src_mean, prm.dst_mean, prm.dst_std = color[0].ptr, color[1].ptr, color[2].ptr
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
