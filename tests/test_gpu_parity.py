"""Parity tests proper: the CUDA library, called through the C ABI, against the CPU oracle on
the same seeded inputs (bit-exact integers, 1e-5 relative on frontier / height / weight),
against the committed golden vectors, and -- at sizes the oracle cannot reach -- through
size-independent properties of the domain.  Run with `-m gpu` on a B200."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from tests import cases

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"
RTOL = 1e-5   # BASELINE.json north_star: float features and link weights within 1e-5 relative


def _seg_host(lib, fg, ct, **kw):
    """ulk_segment_frame_host on numpy buffers."""
    from tests.emu_harness import segment_frame
    return segment_frame(lib, fg, ct, keep_labels=True, **kw)


def _seg_device(lib, fg, ct, **kw):
    """ulk_segment_frame on device buffers (torch owns the memory, non-default stream)."""
    import torch
    from tests.emu_harness import seg_params
    from ultrack_b200 import _lib
    from ultrack_b200.segment import segment_frame_device
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        f = torch.from_numpy(np.ascontiguousarray(fg)).cuda()
        c = torch.from_numpy(np.ascontiguousarray(ct)).cuda()
        s.synchronize()
        h = segment_frame_device(lib, f, c, seg_params(fg_dtype=fg.dtype, **kw), keep_labels=True, stream=s.cuda_stream)
    return h, _lib.nodes_to_host(lib, h, with_labels=True)


@pytest.mark.parametrize("hier", ["area", "dynamics", "volume"])
@pytest.mark.parametrize("kind", cases.KINDS)
@pytest.mark.parametrize("shape", [(56, 56), (10, 24, 24), (160, 160), (24, 48, 48)])
def test_segment_parity(cuda_lib, shape, kind, hier):
    total = 0
    big = shape[-1] > 60 or len(shape) == 3 and shape[0] > 10
    for seed in range(2):
        fg, ct = cases.blob_frame(shape, seed, kind)
        seg = _seg_device if seed else _seg_host
        n = cases.run_and_compare(cuda_lib, seg, fg, ct, len(shape), min_area=40,
                                  max_area=(180 if len(shape) == 2 else 450) * (3 if big else 1), ws_hierarchy=hier,
                                  min_frontier=0.25 if hier == "dynamics" else 0.0)
        total += n or 0
    assert total > 0


@pytest.mark.parametrize("shape", [(160, 160), (20, 12, 272)])
@pytest.mark.parametrize("odd", ["tiny", "negative", "huge", "mixed"])
def test_front_tile_contours_outside_fast_range(cuda_lib, shape, odd):
    """The tile front end computes edge weights with a float expression that is exact for contour
    values that are 0 or in [2^-23, 2^39); a foreground value outside that range raises a flag and
    the edges are recomputed in the general form.  Rows are a multiple of 16 floats (TMA path)."""
    fg, ct = cases.blob_frame(shape, 7, "smooth")
    ct = ct.astype(np.float32)
    rng = np.random.default_rng(11)
    pick = rng.random(shape) < 0.03
    vals = {"tiny": [1e-9, 3e-8, 1e-30], "negative": [-0.25, -1e-3, -0.0], "huge": [1e12, 3e20],
            "mixed": [1e-9, -0.5, 7e11, 3e-8]}[odd]
    ct[pick] = rng.choice(np.array(vals, np.float32), size=int(pick.sum()))
    n = cases.run_and_compare(cuda_lib, _seg_host, fg, ct, len(shape), min_area=30, max_area=2000,
                              ws_hierarchy="area" if odd != "mixed" else "dynamics")
    assert n > 0


@pytest.mark.parametrize("dtype", [np.float32, np.float16, np.uint8])
@pytest.mark.parametrize("thr", [0.3, -0.3, 1e-320, 2.0000000001, 1e39, float("inf")])
def test_threshold_not_representable_in_float32(cuda_lib, dtype, thr):
    """The vectorised threshold kernel compares in float32 against the threshold rounded down to
    float32; the result must equal `double(v) > thr` for thresholds no float32 represents, with
    voxels sitting exactly on either side (C ABI called directly: the Python layer would round thr)."""
    from ultrack_b200 import _lib
    shape = (48, 256)
    mask, ct = cases.blob_frame(shape, 3, "smooth")
    mask = np.asarray(mask) > 0
    rng = np.random.default_rng(0)
    if dtype == np.uint8:
        near = np.array([0, 1, 2, 3, 255], np.uint8)
    else:
        with np.errstate(over="ignore"):
            c = dtype(thr) if np.isfinite(dtype(thr)) else dtype(np.finfo(dtype).max)
            near = np.array([c, np.nextafter(c, dtype(np.inf)), np.nextafter(c, dtype(-np.inf)), 0, np.finfo(dtype).max,
                             -np.finfo(dtype).max, np.inf, -np.inf], dtype)
    above = near[near.astype(np.float64) > thr]
    below = near[~(near.astype(np.float64) > thr)]
    fg = np.where(mask, rng.choice(above, size=shape) if len(above) else below[0],
                  rng.choice(below, size=shape) if len(below) else above[0]).astype(dtype)
    want = int((fg.astype(np.float64) > thr).sum())
    prm = _lib.SegParams(float(thr), 30.0, 4000.0, 4.0, 0.0, 0, 0)
    h = C.POINTER(_lib.FrameNodes)()
    ctc = np.ascontiguousarray(ct, np.float32)
    rc = cuda_lib.ulk_segment_frame_host(fg.ctypes.data_as(C.c_void_p), _lib.dtype_code(fg.dtype), ctc.ctypes.data_as(C.c_void_p),
                                         _lib.dtype_code(ctc.dtype), (C.c_int64 * 3)(1, *shape), C.byref(prm), 0, None, C.byref(h))
    _lib.check(cuda_lib, rc)
    assert h.contents.stats.n_foreground == want
    assert (want > 0) == (len(above) > 0)
    cuda_lib.ulk_frame_nodes_free(h)


@pytest.mark.parametrize("variant,shape,radius", [("T", (20, 40, 40), 6.0), ("T", (128, 128), 7.0),
                                                  ("S", (32, 64, 64), 6.0), ("S", (256, 256), 9.0)])
@pytest.mark.parametrize("min_frontier", [0.0, 0.1])
def test_segment_parity_synthetic_cells(cuda_lib, variant, shape, radius, min_frontier):
    """The bench's own generator (variant S) and the mock-like tie-heavy variant T."""
    from ultrack_b200 import synth
    fg, ct, _ = synth.make_frame(shape, seed=2, variant=variant, radius=radius, fill=0.4)
    n = cases.run_and_compare(cuda_lib, _seg_host, fg, ct, len(shape), min_area=50, max_area=900 if len(shape) == 3 else 400,
                              min_frontier=min_frontier)
    assert n > 0


def test_device_synth_matches_host_recipe(cuda_lib):
    """ulk_synth_frame (bench input) == ultrack_b200.synth.make_frame(variant="S") up to float rounding."""
    import torch
    from ultrack_b200 import _lib, synth
    shape = (8, 48, 40)
    fg = torch.empty(shape, dtype=torch.float32, device="cuda")
    ct = torch.empty(shape, dtype=torch.float32, device="cuda")
    shp = (C.c_int64 * 3)(*shape)
    _lib.check(cuda_lib, cuda_lib.ulk_synth_frame(shp, 3, 11, 6.0, 0.3, 2.0, 1e-3, C.c_void_p(fg.data_ptr()),
                                                  C.c_void_p(ct.data_ptr()), None))
    hfg, hct, _ = synth.make_frame(shape, t=3, seed=11, variant="S", radius=6.0, fill=0.3, motion=2.0, noise=1e-3)
    assert np.allclose(fg.cpu().numpy(), hfg, atol=2e-5)
    assert np.allclose(ct.cpu().numpy(), hct, atol=2e-5)


def test_edge_weights_golden(cuda_lib):
    """Regions cut along the reference's own edge weights: a 2-voxel-wide mask turns the golden
    float32 weights of edges_from_mask into the frontier of a two-basin hierarchy.  Covered
    indirectly: the parity cases above are bit-exact on frontier sums, which are sums of these
    weights; here the dtype classes of the golden file (f32, f64, bool, u8) are segmented."""
    g = np.load(GOLDEN / "edges_cases.npz")
    for name in ("f32_3d", "f64_3d", "bool_3d", "u8_3d", "f32_2d", "bool_2d"):
        mask, img = g[name + "_mask"], g[name + "_img"]
        if mask.shape[0] == 1:
            mask, img = mask[0], img[0]
        cases.run_and_compare(cuda_lib, _seg_host, mask, img, mask.ndim, min_area=4, max_area=60)


def test_iou_golden(cuda_lib):
    """ulk_link_frames reproduces the reference's numba IoU values (tests/golden/iou_cases.npz)."""
    from tests.emu_harness import link_frames
    from ultrack_b200 import frames
    g = np.load(GOLDEN / "iou_cases.npz")
    for k in range(int(g["n"])):
        b1, b2, m1, m2 = g[f"c{k}_b1"], g[f"c{k}_b2"], g[f"c{k}_m1"], g[f"c{k}_m2"]
        d = m1.ndim
        tabs = []
        for b, m in ((b2, m2), (b1, m1)):     # source = node 2, target = node 1 (IoU(target, source))
            bb = np.zeros((1, 6), np.int32)
            bb[0, 3 - d:3] = b[:d]
            bb[0, 6 - d:] = b[d:]
            bb[0, 3:6 - d] = 1
            off, words = frames.pack_masks([m], bb)
            cen = np.zeros((1, 3), np.int32)   # same centroid: distance 0, always a candidate
            tabs.append(frames.upload(cuda_lib, (64, 64, 64), bb, cen, np.array([m.sum()], np.int32), off, words,
                                      np.array([1]), d))
        s, t, w, _ = link_frames(cuda_lib, tabs[0].handle, tabs[1].handle, max_distance=1.0, max_neighbors=1)
        assert len(w) == 1 and w[0] == float(g[f"c{k}_iou"]), (k, w, float(g[f"c{k}_iou"]))


@pytest.mark.parametrize("ndim,scale,dw,with_shift", [(2, None, 0.0, False), (3, None, 0.05, False),
                                                      (3, (2.0, 1.0, 1.0), 0.0, True), (2, None, 0.01, True)])
def test_link_parity(cuda_lib, ndim, scale, dw, with_shift):
    from ultrack_b200 import synth
    shape = (160, 160) if ndim == 2 else (24, 64, 64)
    tabs = []
    for t in range(2):
        fg, ct, _ = synth.make_frame(shape, t=t, seed=11, variant="S", radius=5.0, fill=0.35, motion=2.0)
        tabs.append(_seg_host(cuda_lib, fg, ct, min_area=30, max_area=4000))
    rng = np.random.default_rng(3)
    shift = rng.normal(0, 1.5, size=(len(tabs[1][1]["area"]), 3)) if with_shift else None
    if shift is not None and ndim == 2:
        shift[:, 0] = 0.0
    # only targets whose k-th and (k+1)-th nearest sources are equidistant are left out (< 5 %)
    cases.link_and_compare(cuda_lib, tabs, ndim, 9.0, 2, dw=dw, shift=shift, scale=scale, rtol=RTOL, min_links=30)
    for h, _ in tabs:
        cuda_lib.ulk_frame_nodes_free(h)


def test_labels_to_contours(cuda_lib):
    """utils/edge.py:18-96 restated with scipy grey morphology (skimage find_boundaries, mode="outer")."""
    import scipy.ndimage as ndi
    from ultrack_b200 import labels_to_contours, synth
    sets = []
    for seed in (1, 2, 3):
        frames_ = [synth.make_frame((12, 40, 40), t=t, seed=seed, variant="T", radius=5.0)[2] for t in range(2)]
        sets.append(np.stack(frames_))
    fg, ct = labels_to_contours(sets)

    def find_boundaries_outer(lab):
        nd = lab.ndim
        cross = ndi.generate_binary_structure(nd, 1)
        full = ndi.generate_binary_structure(nd, nd)
        thick = ndi.grey_dilation(lab, footprint=cross) != ndi.grey_erosion(lab, footprint=cross)
        bg = lab == 0
        inv = lab.copy()
        inv[bg] = np.iinfo(lab.dtype).max
        adj = (ndi.grey_dilation(lab, footprint=full) != ndi.grey_erosion(inv, footprint=full)) & ~bg
        return thick & (bg | adj)

    for t in range(2):
        efg = np.zeros(sets[0].shape[1:], bool)
        ect = np.zeros(sets[0].shape[1:], np.float32)
        for lb in sets:
            efg |= lb[t] > 0
            ect += find_boundaries_outer(lb[t])
        ect /= len(sets)
        assert np.array_equal(fg[t], efg)
        assert np.array_equal(ct[t], ect)
    # reference property test (utils/_test/test_edge.py:32-42): foreground covers every label
    for lb in sets:
        assert np.all(fg[lb > 0])


def test_segment_link_api_on_gpu(cuda_lib, tmp_path):
    """segment() + link() through the public API on the GPU; rows against the oracle."""
    from oracle.link_oracle import link_pair
    from oracle.segment_oracle import SegParams, segment_frame as o_seg
    from ultrack_b200 import MainConfig, database, link, segment, synth
    cfg = MainConfig.model_validate({"data": {"working_dir": str(tmp_path)},
                                     "segmentation": {"min_area": 60, "max_area": 2000, "min_frontier": 0.05},
                                     "linking": {"max_distance": 10, "max_neighbors": 3}})
    fg, ct = synth.make_timelapse((16, 64, 64), 3, seed=3, variant="S", radius=5.0, motion=2.0)
    segment(fg, ct, cfg)
    link(cfg)
    with database.connect(cfg.data_config.database_path) as conn:
        rows = conn.execute("select id, area, z, y, x, hier_parent_id from nodes order by id").fetchall()
        links = {(s, t): w for s, t, w in conn.execute("select source_id,target_id,weight from links")}
    prm = SegParams(min_area=60, max_area=2000, min_frontier=0.05)
    ref = [o_seg(fg[t], ct[t], prm, time=t) for t in range(3)]
    exp_rows = [(n["id"], n["area"], n["z"], n["y"], n["x"], n["hier_parent_id"]) for r in ref for n in r["nodes"]]
    assert rows == exp_rows and len(rows) > 20
    exp = {}
    for t in range(2):
        for w, s, d in link_pair(ref[t]["nodes"], ref[t + 1]["nodes"], 10, 3, time=t):
            exp[(int(s), int(d))] = w
    assert set(links) == set(exp)
    for k in exp:
        assert links[k] == pytest.approx(exp[k], rel=RTOL)


def test_full_size_properties(cuda_lib):
    """BASELINE config 3 frame (3D 128x512x512) -- far beyond the oracle's reach -- through
    size-independent properties: determinism (bit-identical reruns), areas = mask popcounts,
    bboxes tight, children nested in parents and overlap pairs = ancestor pairs, regions
    disjoint, identical frames link with exactly one weight-1.0 link per target."""
    import torch
    from tests.emu_harness import link_frames, seg_params
    from ultrack_b200 import _lib
    from ultrack_b200.segment import segment_frame_device
    shape = (128, 512, 512)
    fg = torch.empty(shape, dtype=torch.float32, device="cuda")
    ct = torch.empty(shape, dtype=torch.float32, device="cuda")
    shp = (C.c_int64 * 3)(*shape)
    _lib.check(cuda_lib, cuda_lib.ulk_synth_frame(shp, 0, 5, 8.0, 0.3, 3.0, 1e-3, C.c_void_p(fg.data_ptr()),
                                                  C.c_void_p(ct.data_ptr()), None))
    prm = seg_params(fg_dtype=np.float32, min_area=100, max_area=10000, min_frontier=0.0)
    h0 = segment_frame_device(cuda_lib, fg, ct, prm, keep_labels=True)
    h1 = segment_frame_device(cuda_lib, fg, ct, prm, keep_labels=True)
    a = _lib.nodes_to_host(cuda_lib, h0, with_labels=True)
    b = _lib.nodes_to_host(cuda_lib, h1, with_labels=True)
    for key in ("bbox", "centroid", "area", "frontier", "height", "hier_id", "parent", "hidx", "mask_off", "mask",
                "overlaps", "labels"):
        assert np.array_equal(a[key], b[key]), key
    n = len(a["area"])
    assert n > 1000
    rng = np.random.default_rng(0)
    labels = a["labels"]
    for i in rng.choice(n, 300, replace=False):
        m = _lib.unpack_mask(a, int(i), 3)
        assert m.sum() == a["area"][i]
        assert m.any(axis=(1, 2)).all() and m.any(axis=(0, 2)).all() and m.any(axis=(0, 1)).all()   # tight bbox
        bb = a["bbox"][i]
        crop = labels[bb[0]:bb[3], bb[1]:bb[4], bb[2]:bb[5]]
        assert (crop[m] == a["hier_id"][i]).all()          # the node lies inside its region
        zz, yy, xx = np.nonzero(m)
        cen = np.round(np.array([zz.mean(), yy.mean(), xx.mean()]) + bb[:3]).astype(int)
        assert np.array_equal(cen, a["centroid"][i])
        p = a["parent"][i]
        if p >= 0:
            pb = a["bbox"][p]
            pm = _lib.unpack_mask(a, int(p), 3)
            sub = pm[bb[0] - pb[0]:bb[3] - pb[0], bb[1] - pb[1]:bb[4] - pb[1], bb[2] - pb[2]:bb[5] - pb[2]]
            assert sub.shape == m.shape and (sub[m]).all() and a["area"][p] > a["area"][i]
    # overlaps = all (node, proper emitted ancestor) pairs
    exp = set()
    for i in range(n):
        p = a["parent"][i]
        while p >= 0:
            exp.add((i, int(p)))
            p = a["parent"][p]
    assert {(int(x), int(y)) for x, y in a["overlaps"]} == exp
    # region areas: the root-most emitted node of a region never exceeds the region's voxel count
    counts = np.bincount(labels.ravel())
    roots = np.nonzero(a["parent"] < 0)[0]
    assert (a["area"][roots] <= counts[a["hier_id"][roots]]).all()
    s, t, w, _ = link_frames(cuda_lib, h0, h1, max_distance=15.0, max_neighbors=5)
    ones = np.bincount(t[w == 1.0], minlength=n)
    assert (ones == 1).all()
    assert np.bincount(t, minlength=n).max() <= 5
    cuda_lib.ulk_frame_nodes_free(h0)
    cuda_lib.ulk_frame_nodes_free(h1)


def test_bench_recipe_against_oracle(cuda_lib):
    """The bench's own recipe and parameters (variant S, radius 8, fill 0.3, min_area 100,
    max_area 10000) on the sub-volume the CPU arm times: oversegmentation pass, 128-bit threshold
    path (W % 4 == 0), bottleneck shortcut and boundary-voxel RAG all active, bit-exact vs oracle."""
    from ultrack_b200 import synth
    fg, ct, _ = synth.make_frame((40, 128, 128), t=3, seed=7, variant="S", radius=8.0, fill=0.3, motion=3.0, noise=1e-3)
    n = cases.run_and_compare(cuda_lib, _seg_host, fg, ct, 3, min_area=100, max_area=10000, min_frontier=0.0)
    assert n > 20


def test_shortcuts_equal_exhaustive_forms(cuda_lib, monkeypatch):
    """The data-dependent shortcuts (bottleneck check only where a spanning-tree key exceeds K0,
    components from the foreground bit words, weight bound from K_MinEdge, 128-bit threshold) give
    the same node tables as the exhaustive per-voxel forms they replace, on a frame far beyond the
    oracle's reach (128x256x256)."""
    import torch
    from tests.emu_harness import seg_params
    from ultrack_b200 import _lib
    from ultrack_b200.segment import segment_frame_device
    shape = (128, 256, 256)
    fg = torch.empty(shape, dtype=torch.float32, device="cuda")
    ct = torch.empty(shape, dtype=torch.float32, device="cuda")
    shp = (C.c_int64 * 3)(*shape)
    _lib.check(cuda_lib, cuda_lib.ulk_synth_frame(shp, 1, 11, 8.0, 0.3, 3.0, 1e-3, C.c_void_p(fg.data_ptr()),
                                                  C.c_void_p(ct.data_ptr()), None))
    prm = seg_params(fg_dtype=np.float32, min_area=100, max_area=10000, min_frontier=0.0)

    def run():
        h = segment_frame_device(cuda_lib, fg, ct, prm, keep_labels=True)
        out = _lib.nodes_to_host(cuda_lib, h, with_labels=True)
        cuda_lib.ulk_frame_nodes_free(h)
        return out

    a = run()
    assert 0 < a["stats"]["k3_voxels"] < a["stats"]["n_foreground"]
    for var in ("ULK_K3_ALL", "ULK_CCL_VOXEL", "ULK_CMAX_EARLY", "ULK_THR_SCALAR"):
        monkeypatch.setenv(var, "1")
    b = run()
    assert b["stats"]["k3_voxels"] == b["stats"]["n_foreground"]
    for key in ("bbox", "centroid", "area", "height", "hier_id", "parent", "hidx", "mask_off", "mask", "overlaps", "labels"):
        assert np.array_equal(a[key], b[key]), key
    assert np.allclose(a["frontier"], b["frontier"], rtol=1e-12, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize("ndim,nvecs,fdtype", [(3, 3, np.float32), (3, 2, np.float32), (2, 2, np.float64), (3, 3, np.float16)])
def test_node_shifts_parity(cuda_lib, ndim, nvecs, fdtype):
    """ulk_node_shifts vs the restated add_flow arithmetic (imgproc/flow.py:700-737): field coarser
    than the image, fewer vectors than dimensions, half-way coordinates (round half to even)."""
    import torch
    from oracle.flow_oracle import node_shifts
    from ultrack_b200 import _lib
    shape = (20, 48, 40) if ndim == 3 else (96, 80)
    fshape = (5, 13, 11) if ndim == 3 else (25, 21)
    fg, ct = cases.blob_frame(shape, 3, "smooth")
    h, nodes = _seg_host(cuda_lib, fg, ct, min_area=30, max_area=2000)
    rng = np.random.default_rng(5)
    field = (rng.random((nvecs,) + fshape) - 0.5).astype(fdtype)
    cen = nodes["centroid"][:, 3 - ndim:]
    exp = node_shifts(cen, field, shape)
    dev = torch.from_numpy(field.view(np.uint8).reshape(-1).copy()).cuda()
    fs3 = fshape if ndim == 3 else (1,) + fshape
    out = np.zeros((len(cen), 3), np.float64)
    _lib.check(cuda_lib, cuda_lib.ulk_node_shifts(h, C.c_void_p(dev.data_ptr()), _lib.dtype_code(field.dtype), nvecs,
                                                  (C.c_int64 * 3)(*fs3), out.ctypes.data_as(C.c_void_p), None))
    assert len(cen) > 5 and np.array_equal(out, exp)
    cuda_lib.ulk_frame_nodes_free(h)


def test_add_flow_api_on_gpu(cuda_lib, tmp_path):
    """segment() -> add_flow() -> link() through the public API: the shift columns equal the
    restated reference arithmetic and link() consumes them (imgproc/flow.py:639-751,
    core/linking/processing.py:140-148,211-213)."""
    from oracle.flow_oracle import node_shifts
    from oracle.link_oracle import link_pair
    from oracle.segment_oracle import SegParams, segment_frame as o_seg
    from ultrack_b200 import MainConfig, add_flow, database, link, segment, synth
    cfg = MainConfig.model_validate({"data": {"working_dir": str(tmp_path)},
                                     "segmentation": {"min_area": 60, "max_area": 2000},
                                     "linking": {"max_distance": 10, "max_neighbors": 3}})
    T, shape = 3, (16, 64, 64)
    fg, ct = synth.make_timelapse(shape, T, seed=4, variant="S", radius=5.0, motion=2.0)
    segment(fg, ct, cfg)
    rng = np.random.default_rng(1)
    field = ((rng.random((T, 3, 4, 16, 16)) - 0.5) * 0.2).astype(np.float32)
    add_flow(cfg, field)
    with pytest.raises(ValueError, match="normalized"):
        add_flow(cfg, field * 100)
    with database.connect(cfg.data_config.database_path) as conn:
        rows = conn.execute("select t, z, y, x, z_shift, y_shift, x_shift from nodes order by id").fetchall()
        links = {(s, t): w for s, t, w in conn.execute("select source_id,target_id,weight from links")}
    assert not links
    rows = np.array(rows, float)
    for t in range(T):
        sel = rows[:, 0] == t
        exp = node_shifts(rows[sel, 1:4], field[t], shape)
        assert sel.sum() > 5 and np.array_equal(rows[sel, 4:7], exp) and np.abs(exp).max() > 0
    link(cfg)
    with database.connect(cfg.data_config.database_path) as conn:
        links = {(s, t): w for s, t, w in conn.execute("select source_id,target_id,weight from links")}
    prm = SegParams(min_area=60, max_area=2000)
    ref = [o_seg(fg[t], ct[t], prm, time=t) for t in range(T)]
    exp_links = {}
    for t in range(T - 1):
        sh = rows[rows[:, 0] == t + 1, 4:7]
        for w, s, d in link_pair(ref[t]["nodes"], ref[t + 1]["nodes"], 10, 3, time=t, target_shift=sh):
            exp_links[(int(s), int(d))] = w
    assert set(links) == set(exp_links) and len(links) > 10
    for k in exp_links:
        assert links[k] == pytest.approx(exp_links[k], rel=RTOL)


def test_repeated_runs_stay_exact(cuda_lib):
    """Races show up as rare mismatches: the same frames segmented many times in one process
    (lists shorter than the grid, lanes past the list end, stale scratch from earlier frames)
    must match the oracle every time.  scripts/gpu_hammer.py is the long form."""
    from oracle.segment_oracle import SegParams, segment_frame as o_seg
    from tests.emu_harness import compare_nodes
    todo = []
    for kind, hier in (("smooth", "dynamics"), ("quant", "area")):
        fg, ct = cases.blob_frame((24, 48, 48), 1, kind)
        kw = dict(min_area=40, max_area=1350, ws_hierarchy=hier, min_frontier=0.25 if hier == "dynamics" else 0.0)
        todo.append((fg, ct, kw, o_seg(fg, ct, SegParams(**kw))))
    for rep in range(40):
        for fg, ct, kw, ref in todo:
            h, nodes = _seg_host(cuda_lib, fg, ct, **kw)
            assert np.array_equal(nodes["labels"].reshape(ref["labels"].shape), ref["labels"]), f"labels differ at repetition {rep}"
            if rep % 10 == 0:
                compare_nodes(nodes, ref, 3)
            cuda_lib.ulk_frame_nodes_free(h)


# ---------------------------------------------------------------------------- r02: secondary callers, edge cases, named configs
def _link_tabs_gpu(lib, ndim):
    from ultrack_b200 import synth
    shape = (160, 160) if ndim == 2 else (24, 64, 64)
    tabs = []
    for t in range(2):
        fg, ct, _ = synth.make_frame(shape, t=t, seed=11, variant="S", radius=5.0, fill=0.35, motion=2.0)
        tabs.append(_seg_host(lib, fg, ct, min_area=30, max_area=4000))
    return tabs, shape


@pytest.mark.parametrize("ndim", [2, 3])
@pytest.mark.parametrize("positive", [False, True])
def test_link_intersection_weight(cuda_lib, ndim, positive):
    """weight_func = Node.intersection, edge_must_be_positive (core/match_gt.py:155-168)."""
    tabs, _ = _link_tabs_gpu(cuda_lib, ndim)
    n, ties = cases.link_and_compare(cuda_lib, tabs, ndim, 9.0, 2, dw=0.5, weight_func="intersection", positive=positive,
                                     min_links=30)
    for h, _ in tabs:
        cuda_lib.ulk_frame_nodes_free(h)


@pytest.mark.parametrize("ndim,dtype", [(2, np.float32), (3, np.uint16), (3, np.float64), (3, np.uint8)])
def test_link_color_gate(cuda_lib, ndim, dtype):
    """color_filtering_mask (core/linking/processing.py:52-104): ulk_node_stats + gated ulk_link_frames."""
    tabs, shape = _link_tabs_gpu(cuda_lib, ndim)
    images = cases.color_images(shape, dtype)
    if dtype == np.uint8:
        images = [[(c // 16).astype(np.uint8) for c in fr] for fr in cases.color_images(shape, np.uint16)]
    n, _ = cases.link_and_compare(cuda_lib, tabs, ndim, 9.0, 2, images=images, z_score_threshold=0.8, min_links=20)
    m, _ = cases.link_and_compare(cuda_lib, tabs, ndim, 9.0, 2, min_links=30)
    assert n < m
    for h, _ in tabs:
        cuda_lib.ulk_frame_nodes_free(h)


def test_segment_empty_frame_gpu(cuda_lib):
    fg = np.zeros((16, 16), bool)
    ct = np.zeros((16, 16), np.float32)
    for seg in (_seg_host, _seg_device):
        h, nodes = seg(cuda_lib, fg, ct, min_area=4, max_area=100)
        assert len(nodes["area"]) == 0 and nodes["overlaps"].shape == (0, 2) and not nodes["labels"].any()
        cuda_lib.ulk_frame_nodes_free(h)


def test_segment_full_frame_single_component_gpu(cuda_lib):
    rng = np.random.default_rng(0)
    for shape in ((24, 24), (6, 20, 36)):
        fg = np.ones(shape, bool)
        ct = rng.random(shape).astype(np.float32)
        assert cases.run_and_compare(cuda_lib, _seg_host, fg, ct, len(shape), min_area=10, max_area=120) > 0


def test_region_too_small_raises_like_reference_gpu(cuda_lib):
    """vendored/hierarchy.py:161-162."""
    fg = np.zeros((12, 12), bool)
    fg[2, 2:7] = True          # 1x5 crop: 5 elements < 8
    fg[6:10, 6:10] = True
    ct = np.zeros((12, 12), np.float32)
    with pytest.raises(RuntimeError, match="Region too small. Size of 5 found."):
        _seg_host(cuda_lib, fg, ct, min_area=4, max_area=100)


def test_all_components_removed_gpu(cuda_lib):
    """Every component below the small-object size (hierarchy.py:64-73): no nodes, all-zero labels."""
    rng = np.random.default_rng(3)
    fg = np.zeros((32, 61), bool)
    for _ in range(12):
        y, x = rng.integers(0, 30), rng.integers(0, 58)
        fg[y:y + 2, x:x + 3] = True
    ct = rng.random(fg.shape).astype(np.float32)
    assert cases.run_and_compare(cuda_lib, _seg_host, fg, ct, 2, min_area=100, max_area=1000, min_area_factor=1.0) == 0


def test_config1_mock_timelapse(cuda_lib, tmp_path):
    """BASELINE configs[0]: the reference's conftest timelapse (utils/test_utils.py:106-134: ONE mock
    frame -- bool blobs, bool boundary contours, utils/data.py:15-32 -- repeated 10x at 512x512)
    with make_config_content's parameters (utils/data.py:35-77), through segment() + link().
    Frame parity against the oracle, then the reference's own assertions
    (core/linking/_test/test_link_processing.py:48-63): exactly one weight-1.0 link per target,
    at most max_neighbors links, distances below max_distance, every node linked."""
    from oracle.segment_oracle import SegParams, segment_frame as o_seg
    from ultrack_b200 import MainConfig, database, link, segment, synth
    fg, ct, _ = synth.make_frame((512, 512), seed=42, variant="T", radius=12.0, fill=0.5)
    n = cases.run_and_compare(cuda_lib, _seg_host, fg, ct, 2, **cases.MOCK_CONFIG)
    assert n > 50
    T = 10
    fgs, cts = np.stack([fg] * T), np.stack([ct] * T)
    cfg = MainConfig.model_validate({"data": {"working_dir": str(tmp_path)}, "segmentation": dict(cases.MOCK_CONFIG),
                                     "linking": dict(cases.MOCK_LINK)})
    segment(fgs, cts, cfg)
    link(cfg, scale=(1, 1))
    ref = o_seg(fg, ct, SegParams(**cases.MOCK_CONFIG), time=0)
    with database.connect(cfg.data_config.database_path) as conn:
        nodes = {i: (t, y, x) for i, t, y, x in conn.execute("select id, t, y, x from nodes")}
        links = conn.execute("select source_id, target_id, weight from links").fetchall()
    assert len(nodes) == T * len(ref["nodes"])
    by_target = {}
    for s_, t_, w in links:
        by_target.setdefault(t_, []).append(w)
        (_, ys, xs), (_, yt, xt) = nodes[s_], nodes[t_]
        assert np.hypot(ys - yt, xs - xt) < cases.MOCK_LINK["max_distance"]
    for t_, ws in by_target.items():
        assert len(ws) <= cases.MOCK_LINK["max_neighbors"] and sum(w == 1.0 for w in ws) == 1
    assert set(by_target) == {i for i, (t, _, _) in nodes.items() if t != 0}
    assert {s_ for s_, _, _ in links} == {i for i, (t, _, _) in nodes.items() if t != T - 1}


def test_config2_labels_2048(cuda_lib):
    """BASELINE configs[1] slice: int labels -> labels_to_contours on the device (no host round
    trip) -> segment (area hierarchy) -> link with max_neighbors=5, 2048x2048, two frames, against
    the oracle on the restated contours."""
    import torch
    from ultrack_b200 import labels_to_contours, synth
    from ultrack_b200.capi import seg_params
    from ultrack_b200.segment import segment_frame_device
    from ultrack_b200 import _lib
    labs = np.stack([synth.make_frame((2048, 2048), t=t, seed=3, variant="T", radius=12.0, motion=3.0)[2] for t in range(2)])
    fg_d, ct_d = labels_to_contours(labs.astype(np.int32), device=True)
    assert fg_d.is_cuda and ct_d.is_cuda and fg_d.dtype == torch.bool and ct_d.dtype == torch.float32
    kw = dict(threshold=0.5, min_area=100, max_area=10000)
    tabs = []
    for t in range(2):
        efg, ect = cases.labels_to_contours_ref([labs[t].astype(np.int32)])
        assert np.array_equal(fg_d[t].cpu().numpy(), efg) and np.array_equal(ct_d[t].cpu().numpy(), ect)
        n = cases.run_and_compare(cuda_lib, lambda lib, f, c, **k: _from_device(lib, fg_d[t], ct_d[t], **k), efg, ect, 2, **kw)
        assert n > 1000
        h = segment_frame_device(cuda_lib, fg_d[t], ct_d[t], seg_params(fg_dtype=bool, **kw))
        tabs.append((h, _lib.nodes_to_host(cuda_lib, h)))
    n, ties = cases.link_and_compare(cuda_lib, tabs, 2, 15.0, 5, min_links=1000)
    for h, _ in tabs:
        cuda_lib.ulk_frame_nodes_free(h)


def _from_device(lib, f, c, **kw):
    from ultrack_b200 import _lib
    from ultrack_b200.capi import seg_params
    from ultrack_b200.segment import segment_frame_device
    h = segment_frame_device(lib, f, c, seg_params(fg_dtype=bool, **kw), keep_labels=True)
    return h, _lib.nodes_to_host(lib, h, with_labels=True)


@pytest.mark.parametrize("sigma", [None, 1.0])
def test_config4_ensemble_crop(cuda_lib, sigma):
    """BASELINE configs[3] crop: 4 label hypotheses -> contours in quarters (labels_to_contours),
    dynamics hierarchy + frontier filter, 32x128x128, against the oracle."""
    from ultrack_b200 import labels_to_contours, synth
    base = synth.make_frame((32, 128, 128), seed=9, variant="T", radius=7.0)[2]
    sets = cases.ensemble_label_sets(base, 4, seed=1)
    fg, ct = labels_to_contours([s_[None] for s_ in sets], sigma=sigma)
    efg, ect = cases.labels_to_contours_ref(sets, sigma)
    assert np.array_equal(fg[0], efg)
    if sigma is None:
        assert np.array_equal(ct[0], ect) and set(np.unique(ct[0])) <= {0.0, 0.25, 0.5, 0.75, 1.0}
    else:
        assert np.allclose(ct[0], ect, rtol=1e-5, atol=1e-7)
    n = cases.run_and_compare(cuda_lib, _seg_host, fg[0], ct[0], 3, min_area=100, max_area=5000, ws_hierarchy="dynamics",
                              min_frontier=0.25)
    assert n > 20


def test_bulk_blob_writer_gpu(cuda_lib):
    """ulk_frame_blobs_to_host on the device == the node table, field by field, after unpickling."""
    from ultrack_b200 import _lib, synth
    from ultrack_b200.node import frame_blobs, loads_node
    for shape in ((96, 96), (16, 48, 48)):
        nd = len(shape)
        fg, ct, _ = synth.make_frame(shape, seed=4, variant="S", radius=5.0, fill=0.4)
        h, nodes = _seg_host(cuda_lib, fg, ct, min_area=40, max_area=900)
        buf, off = frame_blobs(cuda_lib, h, nodes, 2, nd, 3_000_000)
        lo = 3 - nd
        assert len(nodes["area"]) > 10
        for i in range(len(nodes["area"])):
            node = loads_node(bytes(buf[off[i]:off[i + 1]]))
            bb = nodes["bbox"][i].astype(np.int64)
            assert np.array_equal(node.bbox, np.concatenate([bb[lo:3], bb[3 + lo:]])) and node.bbox.dtype == np.int64
            assert np.array_equal(node.mask, _lib.unpack_mask(nodes, i, nd)) and node.mask.dtype == bool
            assert np.array_equal(node.centroid, nodes["centroid"][i][lo:])
            assert (node.id, node.time, node.area, node._h_node_index) == (3_000_000 + i + 1, 2, nodes["area"][i], nodes["hidx"][i])
            assert node.frontier == nodes["frontier"][i] and node.height == nodes["height"][i] and node._parent is None
        cuda_lib.ulk_frame_nodes_free(h)


def test_api_with_images_and_properties_gpu(cuda_lib, tmp_path):
    """segment(images=, properties=) + link(images=) through the public API on the GPU."""
    import pickle
    from ultrack_b200 import MainConfig, database, link, segment, synth
    from ultrack_b200.node import loads_node
    cfg = MainConfig.model_validate({"data": {"working_dir": str(tmp_path)},
                                     "segmentation": {"min_area": 60, "max_area": 2000},
                                     "linking": {"max_distance": 10, "max_neighbors": 3, "z_score_threshold": 1.0}})
    shape, T = (12, 48, 48), 3
    fg, ct = synth.make_timelapse(shape, T, seed=3, variant="S", radius=5.0, motion=2.0)
    img = np.stack([cases.color_images(shape, np.float32, seed=t)[0][0] for t in range(T)])
    segment(fg, ct, cfg, images=img, properties=["area", "centroid", "intensity_mean", "intensity_std"])
    link(cfg)
    with database.connect(cfg.data_config.database_path) as conn:
        n_plain = conn.execute("select count(*) from links").fetchone()[0]
    link(cfg, images=[img], overwrite=True)
    with database.connect(cfg.data_config.database_path) as conn:
        n_gated = conn.execute("select count(*) from links").fetchone()[0]
        rows = conn.execute("select t, pickle, features from nodes").fetchall()
    assert 0 < n_gated < n_plain
    for t, blob, feats in rows:
        node = loads_node(blob)
        f = pickle.loads(feats)
        coords = np.stack(np.nonzero(node.mask), axis=1) + np.asarray(node.bbox[:3])
        vals = img[t][tuple(coords.T)].astype(np.float64)
        exp = np.concatenate([[node.area], coords.mean(axis=0), [vals.mean()], [vals.std()]]).astype(np.float32)
        assert np.allclose(f, exp, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", ["plain2d", "dw3d", "scale_shift3d", "shortscale2d", "gt_intersection3d", "positive2d",
                                  "color2d", "color3d_u16"])
def test_link_matches_reference_function_bodies(cuda_lib, name):
    """ulk_link_frames / ulk_node_stats against rows produced by executing the reference's own
    compute_spatial_neighbors + color_filtering_mask source (tests/golden/link_cases.npz)."""
    assert cases.golden_link_check(cuda_lib, name, rtol=RTOL) > 5


def test_bench_recipe_10M_voxels_against_oracle(cuda_lib):
    """The bench recipe and parameters on a 96x320x320 frame (9.8 M voxels, ~2000 nodes, hundreds of
    oversegmented components): the largest input the oracle checks id by id (≈ 20 s of CPU).  The
    frame is generated on the device (ulk_synth_frame) like the bench's own inputs."""
    import torch
    from oracle.compare import compare_nodes
    from oracle.segment_oracle import SegParams, segment_frame as o_seg
    from tests.emu_harness import seg_params
    from ultrack_b200 import _lib
    from ultrack_b200.segment import segment_frame_device
    shape = (96, 320, 320)
    fg = torch.empty(shape, dtype=torch.float32, device="cuda")
    ct = torch.empty(shape, dtype=torch.float32, device="cuda")
    shp = (C.c_int64 * 3)(*shape)
    _lib.check(cuda_lib, cuda_lib.ulk_synth_frame(shp, 1, 7, 8.0, 0.3, 3.0, 1e-3, C.c_void_p(fg.data_ptr()),
                                                  C.c_void_p(ct.data_ptr()), None))
    torch.cuda.synchronize()
    kw = dict(min_area=100, max_area=10000, min_frontier=0.0)
    h = segment_frame_device(cuda_lib, fg, ct, seg_params(fg_dtype=np.float32, **kw), keep_labels=True)
    nodes = _lib.nodes_to_host(cuda_lib, h, with_labels=True)
    ref = o_seg(fg.cpu().numpy(), ct.cpu().numpy(), SegParams(**kw))
    assert len(ref["nodes"]) > 1000 and nodes["stats"]["pass_a_voxels"] > 0
    assert np.array_equal(nodes["labels"], ref["labels"])
    compare_nodes(nodes, ref, 3)
    cuda_lib.ulk_frame_nodes_free(h)
