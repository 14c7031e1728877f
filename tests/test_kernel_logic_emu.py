"""Kernel LOGIC against the oracle without a GPU: the CUDA sources compiled with -DULK_EMU
(tests/emu_harness.py) run thread by thread on the host.  The same cases run against the
real CUDA library in tests/test_gpu_parity.py."""
import numpy as np
import pytest

from tests import cases
from tests.emu_harness import emu_lib, link_frames, segment_frame


def _seg(lib, fg, ct, **kw):
    return segment_frame(lib, fg, ct, keep_labels=True, **kw)


@pytest.mark.parametrize("hier", ["area", "dynamics", "volume"])
@pytest.mark.parametrize("kind", cases.KINDS)
@pytest.mark.parametrize("shape", [(56, 56), (10, 24, 24)])
def test_segment_parity(shape, kind, hier):
    lib = emu_lib()
    total = 0
    for seed in range(2):
        fg, ct = cases.blob_frame(shape, seed, kind)
        n = cases.run_and_compare(lib, _seg, fg, ct, len(shape), min_area=40, max_area=180 if len(shape) == 2 else 450,
                                  ws_hierarchy=hier, min_frontier=0.25 if hier == "dynamics" else 0.0)
        total += n or 0
    assert total > 0


@pytest.mark.parametrize("hier", ["area", "volume"])
@pytest.mark.parametrize("min_frontier", [0.0, 0.3])
def test_segment_parity_mock_like(min_frontier, hier):
    """Variant "T": bool foreground + bool contours, the tie-heavy regime of the reference's
    mock data (utils/data.py:15-32), with the reference's test config ratios."""
    from ultrack_b200 import synth
    lib = emu_lib()
    fg, ct, _ = synth.make_frame((20, 40, 40), seed=2, variant="T", radius=6.0, fill=0.45)
    n = cases.run_and_compare(lib, _seg, fg, ct, 3, min_area=50, max_area=800, min_frontier=min_frontier, ws_hierarchy=hier)
    assert n > 0


def test_segment_no_oversegmentation_needed():
    lib = emu_lib()
    fg, ct = cases.blob_frame((48, 48), 4, "smooth")
    assert cases.run_and_compare(lib, _seg, fg, ct, 2, min_area=40, max_area=1_000_000) > 0


def test_segment_empty_frame():
    lib = emu_lib()
    fg = np.zeros((16, 16), bool)
    ct = np.zeros((16, 16), np.float32)
    h, nodes = _seg(lib, fg, ct, min_area=4, max_area=100)
    assert len(nodes["area"]) == 0 and nodes["overlaps"].shape == (0, 2)
    lib.ulk_frame_nodes_free(h)


def test_segment_full_frame_single_component():
    lib = emu_lib()
    rng = np.random.default_rng(0)
    fg = np.ones((24, 24), bool)
    ct = rng.random((24, 24)).astype(np.float32)
    assert cases.run_and_compare(lib, _seg, fg, ct, 2, min_area=10, max_area=120) > 0


def test_region_too_small_raises_like_reference():
    """vendored/hierarchy.py:161-162."""
    lib = emu_lib()
    fg = np.zeros((12, 12), bool)
    fg[2, 2:7] = True          # 1x5 crop: 5 elements < 8
    fg[6:10, 6:10] = True
    ct = np.zeros((12, 12), np.float32)
    with pytest.raises(RuntimeError, match="Region too small. Size of 5 found."):
        _seg(lib, fg, ct, min_area=4, max_area=100)


def test_golden_edge_inputs_segment_like_the_oracle():
    """The masks / images of the golden edge-weight file (every contour dtype class) as whole
    frames; several contain more than one too-small region, the first in label order raises."""
    from pathlib import Path
    g = np.load(Path(__file__).resolve().parent / "golden" / "edges_cases.npz")
    lib = emu_lib()
    for name in ("f32_3d", "f64_3d", "bool_3d", "u8_3d", "f32_2d", "bool_2d"):
        mask, img = g[name + "_mask"], g[name + "_img"]
        if mask.shape[0] == 1:
            mask, img = mask[0], img[0]
        cases.run_and_compare(lib, _seg, mask, img, mask.ndim, min_area=4, max_area=60)


def _nodes_for_link(nodes, ndim, t):
    from ultrack_b200 import _lib
    out = []
    lo = 3 - ndim
    for i in range(len(nodes["area"])):
        bb = nodes["bbox"][i].astype(np.int64)
        out.append(dict(id=(i + 1) + (t + 1) * 1_000_000, centroid=nodes["centroid"][i][lo:].astype(np.int64),
                        bbox=np.concatenate([bb[lo:3], bb[3 + lo:]]), mask=_lib.unpack_mask(nodes, i, ndim)))
    return out


@pytest.mark.parametrize("ndim,scale,dw,with_shift", [(2, None, 0.0, False), (3, None, 0.05, False),
                                                      (3, (2.0, 1.0, 1.0), 0.0, True), (2, None, 0.01, True)])
def test_link_parity(ndim, scale, dw, with_shift):
    from oracle.link_oracle import link_pair
    from ultrack_b200 import synth
    lib = emu_lib()
    shape = (72, 72) if ndim == 2 else (16, 40, 40)
    frames_ = []
    for t in range(2):
        fg, ct, _ = synth.make_frame(shape, t=t, seed=11, variant="S", radius=5.0, fill=0.35, motion=2.0)
        frames_.append(_seg(lib, fg, ct, min_area=30, max_area=4000))
    (h0, n0), (h1, n1) = frames_
    rng = np.random.default_rng(3)
    shift = rng.normal(0, 1.5, size=(len(n1["area"]), 3)) if with_shift else None
    if shift is not None and ndim == 2:
        shift[:, 0] = 0.0   # the host layer drops the z shift of 2-D frames (link.py)
    max_neighbors, max_distance = 2, 9.0
    s, t, w, npairs = link_frames(lib, h0, h1, max_distance=max_distance, max_neighbors=max_neighbors,
                                  distance_weight=dw, shift=shift, scale=scale)
    src, dst = _nodes_for_link(n0, ndim, 0), _nodes_for_link(n1, ndim, 1)
    ref, cands, dist, _ = link_pair(src, dst, max_distance, max_neighbors, dw, target_shift=shift,
                                    scale=scale, return_candidates=True)
    # k-boundary ties of the KD-tree are heap-order dependent (hard part 5):
    # targets whose candidate list is full are compared only when no tie sits on the boundary
    got = {(int(a) + 1 + 1_000_000, int(b) + 1 + 2_000_000): float(x) for a, b, x in zip(s, t, w)}
    exp = {(int(r[1]), int(r[2])): float(r[0]) for r in ref}
    full = np.isfinite(dist).sum(axis=1) == 2 * max_neighbors
    skip = {dst[i]["id"] for i in np.nonzero(full)[0]}
    got = {k: v for k, v in got.items() if k[1] not in skip}
    exp = {k: v for k, v in exp.items() if k[1] not in skip}
    assert len(exp) > 10
    assert set(got) == set(exp)
    for k in exp:
        assert abs(got[k] - exp[k]) <= 1e-5 * max(abs(exp[k]), 1e-300), (k, got[k], exp[k])
    lib.ulk_frame_nodes_free(h0)
    lib.ulk_frame_nodes_free(h1)


def test_link_identical_frames_have_unit_weight():
    """ultrack/core/linking/_test/test_link_processing.py:48-50: on repeated frames every target
    has exactly one link of weight 1.0 and at most max_neighbors links."""
    from ultrack_b200 import synth
    lib = emu_lib()
    fg, ct, _ = synth.make_frame((64, 64), seed=1, variant="T", radius=6.0)
    h0, n0 = _seg(lib, fg, ct, min_area=20, max_area=600, min_frontier=0.1)
    h1, _ = _seg(lib, fg, ct, min_area=20, max_area=600, min_frontier=0.1)
    s, t, w, _ = link_frames(lib, h0, h1, max_distance=15.0, max_neighbors=10)
    assert len(n0["area"]) > 0
    for j in range(len(n0["area"])):
        wj = w[t == j]
        assert 1 <= len(wj) <= 10 and (wj == 1.0).sum() == 1
    lib.ulk_frame_nodes_free(h0)
    lib.ulk_frame_nodes_free(h1)


def test_link_empty_side():
    import ctypes as C
    from ultrack_b200 import _lib
    lib = emu_lib()
    fg, ct = cases.blob_frame((40, 40), 0, "smooth")
    h0, _ = _seg(lib, fg, ct, min_area=40, max_area=1000)
    h1, n1 = _seg(lib, np.zeros_like(fg), ct, min_area=40, max_area=1000)
    assert len(n1["area"]) == 0
    s, t, w, _ = link_frames(lib, h0, h1)
    assert len(s) == 0
    lib.ulk_frame_nodes_free(h0)
    lib.ulk_frame_nodes_free(h1)
