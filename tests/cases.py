"""Seeded inputs shared by the CPU (emulation) and GPU parity tests."""
from __future__ import annotations

import numpy as np
import scipy.ndimage as ndi

KINDS = ("rand", "smooth", "quant", "bool", "f64", "f16", "u8")


def blob_frame(shape, seed: int, kind: str, fill: float = 0.55):
    """Foreground = thresholded smooth noise (irregular components, holes, border contact);
    contours by `kind`: float32 noise (no ties), smooth, quantised to quarters (ties as in
    BASELINE config 4), bool (the reference mock data), float64, float16, uint8."""
    rng = np.random.default_rng(seed)
    fgf = ndi.gaussian_filter(rng.random(shape), 2.0)
    fg = fgf > np.quantile(fgf, 1.0 - fill)
    sm = ndi.gaussian_filter(rng.random(shape), 1.0)
    if kind == "rand":
        ct = rng.random(shape).astype(np.float32)
    elif kind == "smooth":
        ct = sm.astype(np.float32)
    elif kind == "quant":
        ct = (np.round(sm * 40) / 4 % 1.25).astype(np.float32)
    elif kind == "bool":
        ct = sm > 0.52
    elif kind == "f64":
        ct = sm
    elif kind == "f16":
        ct = sm.astype(np.float16)
    elif kind == "u8":
        ct = np.round(sm * 255).astype(np.uint8)
    else:
        raise ValueError(kind)
    return fg, ct


def run_and_compare(lib, segment_fn, fg, ct, ndim: int, **kw):
    """Segments one frame with `segment_fn(lib, fg, ct, **kw) -> (handle, nodes)` and the
    oracle; asserts the id-keyed parity definition of SURVEY.md section 8a."""
    from oracle.segment_oracle import SegParams, segment_frame as o_seg
    from tests.emu_harness import compare_nodes

    try:
        ref = o_seg(fg, ct, SegParams(**kw))
    except RuntimeError as e:
        ref = e
    try:
        handle, nodes = segment_fn(lib, fg, ct, **kw)
    except RuntimeError as e:
        assert isinstance(ref, RuntimeError), f"library raised {e!r}, oracle did not"
        assert str(e) == str(ref)
        return None
    try:
        assert not isinstance(ref, RuntimeError), f"oracle raised {ref!r}, library did not"
        assert np.array_equal(nodes["labels"].reshape(ref["labels"].shape), ref["labels"]), "region labels differ"
        compare_nodes(nodes, ref, ndim)
    finally:
        lib.ulk_frame_nodes_free(handle)
    return len(ref["nodes"])
