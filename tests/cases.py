"""Seeded inputs shared by the CPU (emulation) and GPU parity tests."""
from __future__ import annotations

from pathlib import Path

import numpy as np
import scipy.ndimage as ndi

GOLDEN_DIR = Path(__file__).resolve().parent / "golden"

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


def nodes_for_link(nodes, ndim, t):
    """Device node table (numpy copy) -> the dict nodes oracle.link_oracle takes."""
    from ultrack_b200 import _lib
    out = []
    lo = 3 - ndim
    for i in range(len(nodes["area"])):
        bb = nodes["bbox"][i].astype(np.int64)
        out.append(dict(id=(i + 1) + (t + 1) * 1_000_000, centroid=nodes["centroid"][i][lo:].astype(np.int64),
                        bbox=np.concatenate([bb[lo:3], bb[3 + lo:]]), mask=_lib.unpack_mask(nodes, i, ndim)))
    return out


def link_and_compare(lib, tabs, ndim, max_distance, max_neighbors, dw=0.0, shift=None, scale=None,
                     weight_func="iou", positive=False, images=None, z_score_threshold=5.0,
                     rtol=1e-5, min_links=10, max_tie_fraction=0.05):
    """`ulk_link_frames` on two segmented frames `tabs = [(handle, nodes), (handle, nodes)]` against
    oracle.link_oracle.link_pair, as keyed (source_id, target_id) sets; weights within `rtol`.
    Targets with a KD-tree k-boundary distance tie are left out (their share is asserted small).
    `images` = (channels of frame t, channels of frame t + 1) turns the colour gate on."""
    from oracle.link_oracle import link_pair
    from ultrack_b200.capi import link_frames, node_intensity
    (h0, n0), (h1, n1) = tabs
    color = frames = None
    if images is not None:
        m0, _ = node_intensity(lib, h0, images[0])
        m1, s1 = node_intensity(lib, h1, images[1])
        color = (m0, m1, s1)
        frames = (np.stack(images[0], axis=-1), np.stack(images[1], axis=-1))
    # the C ABI takes (z, y, x) factors; 2-D frames have z = 0 everywhere
    kscale = None if scale is None else (1.0,) * (3 - len(scale)) + tuple(scale)
    s, t, w, _ = link_frames(lib, h0, h1, max_distance=max_distance, max_neighbors=max_neighbors,
                             distance_weight=dw, shift=shift, scale=kscale, weight_func=weight_func,
                             edge_must_be_positive=positive, color=color, z_score_threshold=z_score_threshold)
    src, dst = nodes_for_link(n0, ndim, 0), nodes_for_link(n1, ndim, 1)
    ref, _, _, _, tie = link_pair(src, dst, max_distance, max_neighbors, dw, target_shift=shift, scale=scale,
                                  return_candidates=True, weight_func=weight_func, edge_must_be_positive=positive,
                                  frames=frames, z_score_threshold=z_score_threshold)
    assert tie.mean() <= max_tie_fraction, f"{tie.sum()} of {len(tie)} targets have a k-boundary tie"
    skip = {dst[i]["id"] for i in np.nonzero(tie)[0]}
    got = {(int(a) + 1 + 1_000_000, int(b) + 1 + 2_000_000): float(x) for a, b, x in zip(s, t, w)}
    exp = {(int(r[1]), int(r[2])): float(r[0]) for r in ref}
    got = {k: v for k, v in got.items() if k[1] not in skip}
    exp = {k: v for k, v in exp.items() if k[1] not in skip}
    assert len(exp) >= min_links, len(exp)
    assert set(got) == set(exp), (len(got), len(exp), sorted(set(got) ^ set(exp))[:5])
    for k in exp:
        assert abs(got[k] - exp[k]) <= rtol * max(abs(exp[k]), 1e-300), (k, got[k], exp[k])
    return len(exp), int(tie.sum())


def color_images(shape, dtype, seed=5):
    """Two image channels for frames t and t + 1: one smooth field per channel plus a small
    frame-to-frame drift and pixel noise, so that a z-score gate of a few sigma removes some
    candidate links and keeps others."""
    rng = np.random.default_rng(seed)
    base = []
    for c in range(2):
        f = ndi.gaussian_filter(rng.random(shape), 3.0)
        base.append((f - f.min()) / (f.max() - f.min()))
    out = []
    for t in range(2):
        chans = []
        for c in range(2):
            d = ndi.gaussian_filter(rng.random(shape), 4.0)
            f = base[c] + 0.3 * t * (d - d.mean()) / d.std() * 0.05 + 0.08 * rng.random(shape)
            if np.issubdtype(np.dtype(dtype), np.integer):
                chans.append(np.round(np.clip(f, 0, None) * 4000).astype(dtype))
            else:
                chans.append(f.astype(dtype))
        out.append(chans)
    return out


# ---------------------------------------------------------------------------- BASELINE.json configs
def find_boundaries_outer(lab: np.ndarray) -> np.ndarray:
    """skimage.segmentation.find_boundaries(lab, mode="outer") restated with scipy grey morphology
    (skimage is not installable here): thick = dilation != erosion over the cross neighbourhood;
    outer keeps the background side plus foreground voxels whose full 3^d neighbourhood holds two
    different non-zero labels."""
    nd = lab.ndim
    cross = ndi.generate_binary_structure(nd, 1)
    full = ndi.generate_binary_structure(nd, nd)
    thick = ndi.grey_dilation(lab, footprint=cross) != ndi.grey_erosion(lab, footprint=cross)
    bg = lab == 0
    inv = lab.copy()
    inv[bg] = np.iinfo(lab.dtype).max
    adj = (ndi.grey_dilation(lab, footprint=full) != ndi.grey_erosion(inv, footprint=full)) & ~bg
    return thick & (bg | adj)


def labels_to_contours_ref(label_sets, sigma=None):
    """utils/edge.py:78-94 for one frame, host restatement."""
    fg = np.zeros(label_sets[0].shape, bool)
    ct = np.zeros(label_sets[0].shape, np.float32)
    for lb in label_sets:
        fg |= lb > 0
        ct += find_boundaries_outer(lb)
    ct /= len(label_sets)
    if sigma is not None:
        ct = ndi.gaussian_filter(ct, sigma)
        ct = ct / ct.max()
    return fg, ct


def ensemble_label_sets(labels: np.ndarray, n_sets: int = 4, seed: int = 0, merge_fraction: float = 0.1):
    """BASELINE config 4: `n_sets` hypotheses = the base labels with independent
    random merges (10 % of the labels take the id of their successor), so the merged contours take
    values in {0, 1/4, 1/2, 3/4, 1}."""
    out = []
    n = int(labels.max())
    for k in range(n_sets):
        rng = np.random.default_rng(seed * 100 + k)
        lut = np.arange(n + 1)
        pick = rng.random(n + 1) < merge_fraction
        pick[0] = False
        lut[pick] = np.minimum(lut[pick] + 1, n)
        out.append(lut[labels].astype(np.int32))
    return out


MOCK_CONFIG = dict(threshold=0.5, min_area=500, max_area=7500, min_frontier=0.1, ws_hierarchy="area")   # utils/data.py:35-77
MOCK_LINK = dict(max_neighbors=10, max_distance=15.0)


def golden_link_check(lib, name: str, rtol: float = 1e-5):
    """`ulk_link_frames` (+ `ulk_node_stats` for the colour gate) against the rows the REFERENCE'S
    own compute_spatial_neighbors produced (tests/golden/link_cases.npz, oracle/make_golden_links.py)."""
    from scipy.spatial import KDTree
    from tests.test_oracle_golden import golden_link_case
    from ultrack_b200 import frames
    from ultrack_b200.capi import link_frames, node_stats
    src, dst, kw, rows, images = golden_link_case(name)
    ndim = len(src[0]["centroid"])
    shape3 = (1,) * (3 - ndim) + tuple(np.load(GOLDEN_DIR / "link_cases.npz")[f"{name}_l0"].shape)

    def table(nodes):
        n = len(nodes)
        bbox3 = np.zeros((n, 6), np.int32)
        cen3 = np.zeros((n, 3), np.int32)
        for i, nd in enumerate(nodes):
            bbox3[i, 3 - ndim:3] = nd["bbox"][:ndim]
            bbox3[i, 6 - ndim:] = nd["bbox"][ndim:]
            bbox3[i, 3:6 - ndim] = 1
            cen3[i, 3 - ndim:] = nd["centroid"]
        off, words = frames.pack_masks([nd["mask"] for nd in nodes], bbox3)
        area = np.array([nd["mask"].sum() for nd in nodes], np.int32)
        return frames.upload(lib, shape3, bbox3, cen3, area, off, words, np.array([nd["id"] for nd in nodes]), ndim)

    ts, td = table(src), table(dst)
    color = None
    if images:
        s0 = node_stats(lib, ts.handle, [im[0] for im in images])
        s1 = node_stats(lib, td.handle, [im[1] for im in images])
        color = (s0["mean"], s1["mean"], s1["std"])
    shift = np.array(kw["target_shift"], np.float64)
    if ndim == 2:
        shift[:, 0] = 0.0
    scale = kw["scale"]
    sc3 = None
    if scale is not None:
        m = min(ndim, len(scale))
        sc3 = [0.0] * (3 - m) + [float(v) for v in list(scale)[-m:]]        # link.link_params: short scales drop leading axes
    s, t, w, _ = link_frames(lib, ts.handle, td.handle, max_distance=kw["max_distance"], max_neighbors=kw["max_neighbors"],
                             distance_weight=kw["distance_weight"], shift=shift, scale=sc3, weight_func=kw["weight_func"],
                             edge_must_be_positive=kw["edge_must_be_positive"], color=color,
                             z_score_threshold=kw.get("z_score_threshold", 5.0))
    got = {(int(ts.ids[a]), int(td.ids[b])): float(x) for a, b, x in zip(s, t, w)}
    exp = {(int(r[1]), int(r[2])): float(r[0]) for r in rows}
    # targets with a k-boundary distance tie (KD-tree heap order) are left out
    k = 2 * kw["max_neighbors"]
    sp = np.asarray([n["centroid"] for n in src])
    tp = np.asarray([n["centroid"] for n in dst], dtype=np.float32) + shift[:, -ndim:]
    if scale is not None:
        m = min(ndim, len(scale))
        sp, tp = sp[..., -m:] * scale[-m:], tp[..., -m:] * scale[-m:]
    skip = set()
    if len(src) > k:
        d1, _ = KDTree(sp).query(tp, k=k + 1, distance_upper_bound=kw["max_distance"])
        tie = np.isfinite(d1[:, k]) & (d1[:, k - 1] == d1[:, k])
        assert tie.mean() <= 0.5, tie.mean()      # the 1-D "short scale" case has many exact ties
        skip = {dst[i]["id"] for i in np.nonzero(tie)[0]}
    got = {key: v for key, v in got.items() if key[1] not in skip}
    exp = {key: v for key, v in exp.items() if key[1] not in skip}
    assert set(got) == set(exp) and len(exp) > 5, (name, len(got), len(exp), sorted(set(got) ^ set(exp))[:6])
    for key in exp:
        assert abs(got[key] - exp[key]) <= rtol * max(abs(exp[key]), 1e-300), (key, got[key], exp[key])
    ts.free()
    td.free()
    return len(exp)
