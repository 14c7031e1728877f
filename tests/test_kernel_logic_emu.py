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


from tests.cases import nodes_for_link as _nodes_for_link  # noqa: E402


def _link_tabs(lib, ndim, shape=None):
    from ultrack_b200 import synth
    shape = shape or ((72, 72) if ndim == 2 else (16, 40, 40))
    tabs = []
    for t in range(2):
        fg, ct, _ = synth.make_frame(shape, t=t, seed=11, variant="S", radius=5.0, fill=0.35, motion=2.0)
        tabs.append(_seg(lib, fg, ct, min_area=30, max_area=4000))
    return tabs, shape


@pytest.mark.parametrize("ndim,scale,dw,with_shift", [(2, None, 0.0, False), (3, None, 0.05, False),
                                                      (3, (2.0, 1.0, 1.0), 0.0, True), (2, None, 0.01, True)])
def test_link_parity(ndim, scale, dw, with_shift):
    lib = emu_lib()
    tabs, _ = _link_tabs(lib, ndim)
    rng = np.random.default_rng(3)
    shift = rng.normal(0, 1.5, size=(len(tabs[1][1]["area"]), 3)) if with_shift else None
    if shift is not None and ndim == 2:
        shift[:, 0] = 0.0   # the host layer drops the z shift of 2-D frames (link.py)
    cases.link_and_compare(lib, tabs, ndim, 9.0, 2, dw=dw, shift=shift, scale=scale)
    for h, _ in tabs:
        lib.ulk_frame_nodes_free(h)


@pytest.mark.parametrize("ndim", [2, 3])
@pytest.mark.parametrize("positive", [False, True])
def test_link_intersection_weight(ndim, positive):
    """The GT matcher's call of compute_spatial_neighbors (core/match_gt.py:155-168):
    weight = Node.intersection, edge_must_be_positive."""
    lib = emu_lib()
    tabs, _ = _link_tabs(lib, ndim)
    cases.link_and_compare(lib, tabs, ndim, 9.0, 2, dw=0.5, weight_func="intersection", positive=positive)
    for h, _ in tabs:
        lib.ulk_frame_nodes_free(h)


@pytest.mark.parametrize("ndim,dtype", [(2, np.float32), (3, np.uint16), (3, np.float64)])
def test_link_color_gate(ndim, dtype):
    """color_filtering_mask (core/linking/processing.py:52-104) with two image channels."""
    lib = emu_lib()
    tabs, shape = _link_tabs(lib, ndim)
    images = cases.color_images(shape, dtype)
    n, _ = cases.link_and_compare(lib, tabs, ndim, 9.0, 2, images=images, z_score_threshold=0.8)
    m, _ = cases.link_and_compare(lib, tabs, ndim, 9.0, 2)
    assert n < m            # the gate removed links
    for h, _ in tabs:
        lib.ulk_frame_nodes_free(h)


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


@pytest.mark.parametrize("shape,sigma", [((40, 36), 1.5), ((9, 20, 24), (0.5, 1.0, 2.0)), ((6, 18, 18), 3.0)])
def test_smooth_contours_matches_scipy(shape, sigma):
    """ulk_smooth_contours == scipy.ndimage.gaussian_filter + max-normalise (utils/edge.py:89-91)."""
    import ctypes as C
    import scipy.ndimage as ndi
    from ultrack_b200 import _lib
    from ultrack_b200.edge import _sigma3
    lib = emu_lib()
    rng = np.random.default_rng(0)
    ct = (rng.random(shape) > 0.8).astype(np.float32) * rng.choice([0.25, 0.5, 0.75, 1.0], size=shape).astype(np.float32)
    exp = ndi.gaussian_filter(ct, sigma)
    exp = exp / exp.max()
    got = np.ascontiguousarray(ct.copy())
    shape3 = shape if len(shape) == 3 else (1,) + shape
    _lib.check(lib, lib.ulk_smooth_contours(got.ctypes.data_as(C.c_void_p), (C.c_int64 * 3)(*shape3),
                                            (C.c_double * 3)(*_sigma3(sigma, len(shape))), None))
    assert np.allclose(got, exp, rtol=1e-5, atol=1e-7)
    assert np.abs(got - exp).max() <= 2e-7


@pytest.mark.parametrize("shape,sigma", [((40, 36), 1.5), ((9, 20, 24), (0.5, 1.0, 2.0)), ((6, 18, 18), 3.0), ((12, 10), (0.0, 2.0))])
def test_smooth_contours_static_equals_allocating_entry(shape, sigma):
    """ulk_smooth_contours_static (fixed scratch, taps in the kernel parameters, device-side maximum)
    gives the bits of ulk_smooth_contours, twice in a row on the shared scratch; limits are refused."""
    import ctypes as C
    from ultrack_b200 import _lib
    from ultrack_b200.edge import _sigma3
    lib = emu_lib()
    shape3 = shape if len(shape) == 3 else (1,) + shape
    shp, sg = (C.c_int64 * 3)(*shape3), (C.c_double * 3)(*_sigma3(sigma, len(shape)))
    for seed in (0, 1):
        rng = np.random.default_rng(seed)
        ct = (rng.random(shape) > 0.8).astype(np.float32) * rng.choice([0.25, 0.5, 0.75, 1.0], size=shape).astype(np.float32)
        ref, got = ct.copy(), ct.copy()
        _lib.check(lib, lib.ulk_smooth_contours(ref.ctypes.data_as(C.c_void_p), shp, sg, None))
        _lib.check(lib, lib.ulk_smooth_contours_static(got.ctypes.data_as(C.c_void_p), shp, sg, None))
        assert np.array_equal(ref.view(np.uint32), got.view(np.uint32))
    big = (C.c_int64 * 3)(1, 4096, 2048)
    assert lib.ulk_smooth_contours_static(got.ctypes.data_as(C.c_void_p), big, sg, None) == -12
    assert b"static scratch limit" in lib.ulk_last_error()
    assert lib.ulk_smooth_contours_static(got.ctypes.data_as(C.c_void_p), shp, (C.c_double * 3)(0.0, 40.0, 1.0), None) == -12
    assert lib.ulk_smooth_contours_static_max_voxels() == 4194304


@pytest.mark.parametrize("name", ["plain2d", "dw3d", "scale_shift3d", "shortscale2d", "gt_intersection3d", "positive2d",
                                  "color2d", "color3d_u16"])
def test_link_matches_reference_function_bodies(name):
    """The link kernels against rows produced by the reference's own compute_spatial_neighbors source."""
    assert cases.golden_link_check(emu_lib(), name) > 5


@pytest.mark.parametrize("seed", range(24))
def test_segment_parity_fuzz(seed):
    """Seeded random cases over shape, fill, contour kind, hierarchy attribute and every config
    field the path reads (min_area, min_area_factor, max_area, min_frontier, threshold)."""
    rng = np.random.default_rng(1000 + seed)
    nd = int(rng.integers(2, 4))
    shape = tuple(int(v) for v in (rng.integers(20, 64, 2) if nd == 2 else rng.integers(6, 26, 3)))
    kind = cases.KINDS[int(rng.integers(0, len(cases.KINDS)))]
    fg, ct = cases.blob_frame(shape, seed, kind, fill=float(rng.uniform(0.15, 0.8)))
    hier = ("area", "dynamics", "volume")[int(rng.integers(0, 3))]
    min_area = int(rng.integers(4, 60))
    kw = dict(min_area=min_area, max_area=int(min_area * rng.uniform(1.5, 30)), ws_hierarchy=hier,
              min_area_factor=float(rng.choice([1.0, 2.0, 4.0, 8.0])), min_frontier=float(rng.choice([0.0, 0.0, 0.1, 0.3])))
    cases.run_and_compare(emu_lib(), _seg, fg, ct, nd, **kw)
