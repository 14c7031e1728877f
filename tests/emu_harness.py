"""Host-emulation harness for the CPU test tier (TEST INFRASTRUCTURE ONLY).

Builds tests/emu/libulk_emu.so from the SAME kernel sources as the CUDA library with
-DULK_EMU (kernel functors executed thread by thread on the host) so that the kernel
logic -- basins, join keys, hierarchy ordering, link selection -- is checked against the
oracle without a GPU.  The package itself never loads this library.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import numpy as np

from ultrack_b200 import _lib
from ultrack_b200.csrc.build import build_emu

_EMU = None


def emu_lib():
    global _EMU
    if _EMU is None:
        _EMU = _lib.bind(C.CDLL(str(build_emu())))
    return _EMU


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
    entry = lib.ulk_segment_frame if lib is _EMU else lib.ulk_segment_frame_host
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


def compare_nodes(nodes: Dict[str, np.ndarray], ref: dict, ndim: int, t: int = 0,
                  max_segments: int = 1_000_000, rtol: float = 1e-5) -> None:
    """id-keyed comparison of a device node table with the oracle's output
    (SURVEY.md section 8a parity definition)."""
    rn = ref["nodes"]
    assert len(rn) == len(nodes["area"]), f"node count {len(nodes['area'])} != oracle {len(rn)}"
    for i, r in enumerate(rn):
        assert r["t_node_id"] == i + 1
        assert r["id"] == (i + 1) + (t + 1) * max_segments
        assert r["t_hier_id"] == nodes["hier_id"][i], (i, r["t_hier_id"], nodes["hier_id"][i])
        assert r["area"] == nodes["area"][i], (i, r["area"], nodes["area"][i])
        bb = nodes["bbox"][i]
        rb = r["bbox"]
        if ndim == 2:
            got = np.array([bb[1], bb[2], bb[4], bb[5]])
            assert bb[0] == 0 and bb[3] == 1
        else:
            got = bb
        assert np.array_equal(got, rb), (i, got, rb)
        c = nodes["centroid"][i]
        assert (c[0] if ndim == 3 else 0, c[1], c[2]) == (r["z"], r["y"], r["x"]), (i, c, r["z"], r["y"], r["x"])
        m = _lib.unpack_mask(nodes, i, ndim)
        assert m.shape == r["mask"].shape and np.array_equal(m, r["mask"]), f"mask of node {i} differs"
        assert r["h_node_index"] == nodes["hidx"][i], (i, r["h_node_index"], nodes["hidx"][i])
        for key in ("frontier", "height"):
            a, b = float(nodes[key][i]), float(r[key])
            assert a == b or abs(a - b) <= rtol * max(abs(a), abs(b)), (i, key, a, b)
        par = nodes["parent"][i]
        pid = -1 if par < 0 else (par + 1) + (t + 1) * max_segments
        assert pid == r["hier_parent_id"], (i, pid, r["hier_parent_id"])
    ov = {((a + 1) + (t + 1) * max_segments, (b + 1) + (t + 1) * max_segments) for a, b in nodes["overlaps"]}
    assert ov == set(ref["overlaps"]), "overlap pair sets differ"
