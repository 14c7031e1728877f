"""ctypes binding of the C ABI declared in include/ultrack_b200.h.

The product path loads ONLY ultrack_b200/csrc/libultrack_b200.so (hand-written sm_100a
CUDA) and raises if it is missing or no CUDA device is visible: there is no CPU fallback.
`bind()` is also used by tests/emu to bind the host-emulation test harness.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Dict, Optional

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "csrc" / "libultrack_b200.so"

DTYPE_CODES = {
    np.dtype(np.bool_): 0, np.dtype(np.uint8): 0, np.dtype(np.int8): 1, np.dtype(np.uint16): 2,
    np.dtype(np.int16): 3, np.dtype(np.float16): 4, np.dtype(np.float32): 5, np.dtype(np.float64): 6,
    np.dtype(np.int32): 7, np.dtype(np.uint32): 8, np.dtype(np.int64): 9, np.dtype(np.uint64): 10,
}
WS_HIERARCHY_CODES = {"area": 0, "dynamics": 1, "volume": 2}
WEIGHT_KIND_CODES = {"iou": 0, "intersection": 1}


class SegParams(C.Structure):
    _fields_ = [("threshold", C.c_double), ("min_area", C.c_double), ("max_area", C.c_double),
                ("min_area_factor", C.c_double), ("min_frontier", C.c_double),
                ("ws_hierarchy", C.c_int32), ("reserved", C.c_int32)]


class SegStats(C.Structure):
    _fields_ = [("n_basins", C.c_int64), ("n_rag_edges", C.c_int64), ("plateau_voxels", C.c_int64),
                ("flow_rounds", C.c_int32), ("k3_rounds", C.c_int32),
                ("n_foreground", C.c_int64), ("pass_a_voxels", C.c_int64),
                ("k3_voxels", C.c_int64)]


class FrameNodes(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("n_overlaps", C.c_int32), ("n_hier", C.c_int32),
                ("fix_shift", C.c_int32), ("shape", C.c_int32 * 3), ("pad0", C.c_int32),
                ("mask_words", C.c_int64),
                ("bbox", C.c_void_p), ("centroid", C.c_void_p), ("area", C.c_void_p),
                ("frontier", C.c_void_p), ("height", C.c_void_p), ("hier_id", C.c_void_p),
                ("parent", C.c_void_p), ("hidx", C.c_void_p), ("mask_off", C.c_void_p),
                ("mask", C.c_void_p), ("overlaps", C.c_void_p), ("labels", C.c_void_p),
                ("stats", SegStats)]


class LinkParams(C.Structure):
    _fields_ = [("max_distance", C.c_double), ("distance_weight", C.c_double),
                ("max_neighbors", C.c_int32), ("has_scale", C.c_int32), ("scale", C.c_double * 3),
                ("weight_kind", C.c_int32), ("edge_must_be_positive", C.c_int32),
                ("n_channels", C.c_int32), ("pad0", C.c_int32), ("z_score_threshold", C.c_double),
                ("src_mean", C.c_void_p), ("dst_mean", C.c_void_p), ("dst_std", C.c_void_p)]


class BlobSpec(C.Structure):
    _fields_ = [(k, C.c_int32) for k in (
        "head_len", "tail_len", "ndim", "payload_start", "payload_tail",
        "off_area", "off_frontier", "off_height", "off_hidx", "off_bbox", "off_chunk_len",
        "off_blosc_nbytes", "off_blosc_blocksize", "off_blosc_cbytes", "off_shape", "off_raw_len",
        "toff_id", "toff_time", "toff_centroid")]


class Links(C.Structure):
    _fields_ = [("n_links", C.c_int64), ("source", C.c_void_p), ("target", C.c_void_p),
                ("weight", C.c_void_p), ("n_pairs", C.c_int64)]


EXPORTS = [
    "ulk_last_error", "ulk_device_count", "ulk_set_device", "ulk_launch_count", "ulk_workspace_stats",
    "ulk_profile_enable", "ulk_profile_reset", "ulk_profile_report", "ulk_segment_frame",
    "ulk_segment_frame_host", "ulk_frame_nodes_free", "ulk_frame_nodes_to_host",
    "ulk_frame_nodes_from_host", "ulk_frame_nodes_to_device", "ulk_frame_nodes_from_device", "ulk_link_frames", "ulk_links_to_host", "ulk_links_free",
    "ulk_labels_to_contours", "ulk_node_shifts", "ulk_synth_frame", "ulk_node_stats", "ulk_frame_blobs_to_host", "ulk_smooth_contours",
    "ulk_upload_begin", "ulk_segment_frame_uploaded", "ulk_upload_free",
    "ulk_smooth_contours_static", "ulk_smooth_contours_static_max_voxels",
]


def bind(lib: C.CDLL) -> C.CDLL:
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    PF, PL = C.POINTER(FrameNodes), C.POINTER(Links)
    lib.ulk_last_error.restype = C.c_char_p
    lib.ulk_last_error.argtypes = []
    lib.ulk_device_count.restype = C.c_int
    lib.ulk_set_device.argtypes = [C.c_int]
    lib.ulk_launch_count.restype = i64
    lib.ulk_workspace_stats.argtypes = [C.POINTER(i64)]
    lib.ulk_workspace_stats.restype = None
    lib.ulk_profile_enable.argtypes = [C.c_int]
    lib.ulk_profile_enable.restype = None
    lib.ulk_profile_reset.restype = None
    lib.ulk_profile_report.argtypes = [C.c_char_p, i64]
    seg_args = [vp, C.c_int, vp, C.c_int, C.POINTER(i64), C.POINTER(SegParams), C.c_int, vp, C.POINTER(PF)]
    lib.ulk_segment_frame.argtypes = seg_args
    lib.ulk_segment_frame_host.argtypes = seg_args
    lib.ulk_frame_nodes_free.argtypes = [PF]
    lib.ulk_frame_nodes_free.restype = None
    lib.ulk_upload_begin.argtypes = [vp, C.c_int, vp, C.c_int, C.POINTER(i64), C.POINTER(vp)]
    lib.ulk_segment_frame_uploaded.argtypes = [vp, C.POINTER(SegParams), C.c_int, vp, C.POINTER(PF)]
    lib.ulk_upload_free.argtypes = [vp]
    lib.ulk_upload_free.restype = None
    lib.ulk_frame_nodes_to_host.argtypes = [PF] + [vp] * 12 + [vp]
    lib.ulk_frame_nodes_from_host.argtypes = [i32, C.POINTER(i32), vp, vp, vp, vp, vp, vp, C.POINTER(PF)]
    lib.ulk_frame_nodes_to_device.argtypes = [PF, vp, vp, vp, vp, vp, vp]
    lib.ulk_frame_nodes_from_device.argtypes = [i32, C.POINTER(i32), vp, vp, vp, vp, i64, vp, vp, C.POINTER(PF)]
    lib.ulk_link_frames.argtypes = [PF, PF, vp, C.POINTER(LinkParams), vp, C.POINTER(PL)]
    lib.ulk_links_to_host.argtypes = [PL, vp, vp, vp, vp]
    lib.ulk_frame_blobs_to_host.argtypes = [PF, vp, vp, C.POINTER(BlobSpec), i64, i32, vp, vp, vp]
    lib.ulk_smooth_contours.argtypes = [vp, C.POINTER(i64), C.POINTER(dbl), vp]
    lib.ulk_smooth_contours_static.argtypes = [vp, C.POINTER(i64), C.POINTER(dbl), vp]
    lib.ulk_smooth_contours_static_max_voxels.argtypes = []
    lib.ulk_smooth_contours_static_max_voxels.restype = C.c_int64
    lib.ulk_node_stats.argtypes = [PF, vp, C.c_int, i32, i32, vp, vp, vp, vp, vp, vp]
    lib.ulk_links_free.argtypes = [PL]
    lib.ulk_links_free.restype = None
    lib.ulk_labels_to_contours.argtypes = [vp, i32, C.POINTER(i64), vp, vp, vp]
    lib.ulk_node_shifts.argtypes = [PF, vp, C.c_int, i32, C.POINTER(i64), vp, vp]
    lib.ulk_synth_frame.argtypes = [C.POINTER(i64), i32, C.c_uint64, dbl, dbl, dbl, dbl, vp, vp, vp]
    return lib


_LIB: Optional[C.CDLL] = None


def get_lib() -> C.CDLL:
    """The CUDA library.  Raises (never falls back) when it cannot run."""
    global _LIB
    if _LIB is None:
        if not LIB_PATH.exists():
            # a fresh checkout has no binary (*.so is not tracked): build it once when nvcc is here
            try:
                from ultrack_b200.csrc.build import build
                build()
            except Exception as e:  # noqa: BLE001
                raise RuntimeError(
                    f"{LIB_PATH} is missing and could not be built ({e}): run `python -m ultrack_b200.csrc.build` "
                    "(nvcc, sm_100a). ultrack_b200 has no CPU fallback.") from e
        lib = bind(C.CDLL(str(LIB_PATH)))
        if lib.ulk_device_count() <= 0:
            raise RuntimeError("ultrack_b200 needs a CUDA device (B200, sm_100a); none is visible "
                               "and there is no CPU fallback.")
        _LIB = lib
    return _LIB


def profile_report(lib: C.CDLL) -> Dict[str, tuple]:
    """{kernel name: (launches, total ms)} accumulated since the last ulk_profile_reset."""
    need = lib.ulk_profile_report(None, 0)
    buf = C.create_string_buffer(int(need) + 16)
    lib.ulk_profile_report(buf, len(buf))
    out = {}
    for line in buf.value.decode("utf-8", "replace").splitlines():
        name, cnt, ms = line.rsplit("\t", 2)
        out[name] = (int(cnt), float(ms))
    return out


class UlkError(RuntimeError):
    pass


def check(lib: C.CDLL, rc: int) -> None:
    if rc == 0:
        return
    msg = (lib.ulk_last_error() or b"").decode("utf-8", "replace")
    if rc == -10:
        # vendored/hierarchy.py:161-162 raises RuntimeError for tiny regions
        raise RuntimeError(msg)
    if rc == -1:
        raise ValueError(msg)
    raise UlkError(f"ultrack_b200 native call failed ({rc}): {msg}")


def dtype_code(dt) -> int:
    dt = np.dtype(dt)
    if dt not in DTYPE_CODES:
        raise TypeError(f"unsupported array dtype {dt}")
    return DTYPE_CODES[dt]


_PINNED: Dict[str, "np.ndarray"] = {}


def _host_array(key: str, shape, dtype, pinned: bool) -> np.ndarray:
    """Output buffer of `nodes_to_host`: fresh numpy memory, or (pinned=True) a view of a cached
    page-locked buffer that is reused by the next call -- copies then run at PCIe speed."""
    if not pinned:
        return np.zeros(shape, dtype)
    import torch
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    buf = _PINNED.get(key)
    if buf is None or buf.nbytes < n:
        buf = torch.empty(max(n + n // 4, 4096), dtype=torch.uint8, pin_memory=True).numpy()
        _PINNED[key] = buf
    return buf[:n].view(dtype).reshape(shape)


def nodes_to_host(lib: C.CDLL, handle, stream: int = 0, with_labels: bool = False,
                  with_masks: bool = True, pinned: bool = False) -> Dict[str, np.ndarray]:
    """Copies a device node table into numpy arrays (`with_masks=False`: everything but the
    bit-packed masks, which `segment()` does not need on the host -- the blobs are written by
    `ulk_frame_blobs_to_host`)."""
    f = handle.contents
    n, m = int(f.n_nodes), int(f.n_overlaps)
    z = lambda key, shape, dt: _host_array(key, shape, dt, pinned)  # noqa: E731
    out = dict(
        bbox=z("bbox", (n, 6), np.int32), centroid=z("centroid", (n, 3), np.int32), area=z("area", (n,), np.int32),
        frontier=z("frontier", (n,), np.float64), height=z("height", (n,), np.float64), hier_id=z("hier_id", (n,), np.int32),
        parent=z("parent", (n,), np.int32), hidx=z("hidx", (n,), np.int32), mask_off=z("mask_off", (n + 1,), np.int64),
        mask=z("mask", (int(f.mask_words) if with_masks else 0,), np.uint32), overlaps=z("overlaps", (m, 2), np.int32),
    )
    labels = None
    if with_labels and f.labels:
        labels = np.zeros(tuple(f.shape), np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None  # noqa: E731
    check(lib, lib.ulk_frame_nodes_to_host(
        handle, p(out["bbox"]), p(out["centroid"]), p(out["area"]), p(out["frontier"]), p(out["height"]),
        p(out["hier_id"]), p(out["parent"]), p(out["hidx"]), p(out["mask_off"]), p(out["mask"]) if with_masks else None,
        p(out["overlaps"]), p(labels), C.c_void_p(stream)))
    out["n_hier"] = int(f.n_hier)
    out["shape"] = tuple(int(s) for s in f.shape)
    out["stats"] = {k: int(getattr(f.stats, k)) for k, _ in SegStats._fields_}
    if labels is not None:
        out["labels"] = labels
    return out


def unpack_mask(nodes: Dict[str, np.ndarray], i: int, ndim: int = 3) -> np.ndarray:
    """Bool mask of node i over its bbox (the reference's Node.mask)."""
    bb = nodes["bbox"][i]
    d, h, w = int(bb[3] - bb[0]), int(bb[4] - bb[1]), int(bb[5] - bb[2])
    wpr = (w + 31) // 32
    words = nodes["mask"][int(nodes["mask_off"][i]):int(nodes["mask_off"][i + 1])]
    bits = np.unpackbits(words.view(np.uint8), bitorder="little").reshape(d, h, wpr * 32)[:, :, :w]
    m = bits.astype(bool)
    return m[0] if ndim == 2 else m
