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
    from oracle.link_oracle import link_pair
    from tests.emu_harness import link_frames
    from tests.test_kernel_logic_emu import _nodes_for_link
    from ultrack_b200 import synth
    shape = (160, 160) if ndim == 2 else (24, 64, 64)
    tabs = []
    for t in range(2):
        fg, ct, _ = synth.make_frame(shape, t=t, seed=11, variant="S", radius=5.0, fill=0.35, motion=2.0)
        tabs.append(_seg_host(cuda_lib, fg, ct, min_area=30, max_area=4000))
    (h0, n0), (h1, n1) = tabs
    rng = np.random.default_rng(3)
    shift = rng.normal(0, 1.5, size=(len(n1["area"]), 3)) if with_shift else None
    if shift is not None and ndim == 2:
        shift[:, 0] = 0.0
    max_neighbors, max_distance = 2, 9.0
    s, t, w, _ = link_frames(cuda_lib, h0, h1, max_distance=max_distance, max_neighbors=max_neighbors,
                             distance_weight=dw, shift=shift, scale=scale)
    src, dst = _nodes_for_link(n0, ndim, 0), _nodes_for_link(n1, ndim, 1)
    ref, _, dist, _ = link_pair(src, dst, max_distance, max_neighbors, dw, target_shift=shift, scale=scale,
                                return_candidates=True)
    full = np.isfinite(dist).sum(axis=1) == 2 * max_neighbors   # k-boundary ties are KD-tree internals
    skip = {dst[i]["id"] for i in np.nonzero(full)[0]}
    got = {(int(a) + 1 + 1_000_000, int(b) + 1 + 2_000_000): float(x) for a, b, x in zip(s, t, w)}
    exp = {(int(r[1]), int(r[2])): float(r[0]) for r in ref}
    got = {k: v for k, v in got.items() if k[1] not in skip}
    exp = {k: v for k, v in exp.items() if k[1] not in skip}
    assert set(got) == set(exp) and len(exp) > 30
    for k in exp:
        assert abs(got[k] - exp[k]) <= RTOL * max(abs(exp[k]), 1e-300)
    cuda_lib.ulk_frame_nodes_free(h0)
    cuda_lib.ulk_frame_nodes_free(h1)


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
