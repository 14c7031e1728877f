"""segment() / link() end to end on the host-emulation harness: SQL schema and rows against
the oracle, the reference's own invariants (ultrack/core/segmentation/_test/
test_segment_processing.py:15-107, ultrack/core/linking/_test/test_link_processing.py:16-115),
blob format, config mirror, error behaviour."""
import pickle
import sqlite3
import warnings

import numpy as np
import pytest

from oracle.link_oracle import iou, link_pair
from oracle.segment_oracle import SegParams, segment_frame as o_seg
from ultrack_b200 import MainConfig, database, link, segment, synth
from ultrack_b200.config import load_config, save_config, ws_hierarchy_name
from ultrack_b200.node import dumps_node, loads_node, pack_array, unpack_array


def _config(tmp_path, **over):
    d = {"data": {"working_dir": str(tmp_path)},
         "segmentation": {"min_area": 60, "max_area": 2000, "min_frontier": 0.0, "threshold": 0.5},
         "linking": {"max_distance": 10, "max_neighbors": 3}}
    for k, v in over.items():
        sec, key = k.split("__")
        d[sec][key] = v
    return MainConfig.model_validate(d)


def _rows(cfg, sql):
    with database.connect(cfg.data_config.database_path) as conn:
        return conn.execute(sql).fetchall()


@pytest.mark.parametrize("shape", [(96, 96), (12, 40, 40)])
def test_segment_link_rows_match_oracle(tmp_path, emu, shape):
    cfg = _config(tmp_path)
    T = 3
    fg, ct = synth.make_timelapse(shape, T, seed=3, variant="S", radius=7.0 if len(shape) == 2 else 5.0, motion=2.0)
    segment(fg, ct, cfg)
    link(cfg)
    assert cfg.data_config.metadata["shape"] == [T, *shape]
    rows = _rows(cfg, "select t,id,t_node_id,t_hier_id,z,y,x,area,frontier,height,hier_parent_id,pickle,"
                      "parent_id,selected,node_prob,z_shift,segm_annot,features from nodes order by id")
    prm = SegParams(threshold=0.5, min_area=60, max_area=2000)
    ref = [o_seg(fg[t], ct[t], prm, time=t) for t in range(T)]
    allref = [n for r in ref for n in r["nodes"]]
    assert len(allref) == len(rows) > 10
    for r, n in zip(rows, allref):
        assert r[:8] == (n["t"], n["id"], n["t_node_id"], n["t_hier_id"], n["z"], n["y"], n["x"], n["area"])
        assert r[8] == pytest.approx(n["frontier"], rel=1e-5) and r[9] == pytest.approx(n["height"], rel=1e-5)
        assert r[10] == n["hier_parent_id"]
        assert r[12:] == (-1, 0, -1.0, 0.0, "UNKNOWN", None)
        node = loads_node(r[11])
        assert np.array_equal(node.mask, n["mask"]) and node.mask.dtype == bool
        assert np.array_equal(node.bbox, n["bbox"]) and np.array_equal(node.centroid, n["centroid"])
        assert (node.id, node.time, node.area, node._h_node_index) == (n["id"], n["t"], n["area"], n["h_node_index"])
    ov = set(_rows(cfg, "select node_id, ancestor_id from overlaps"))
    assert ov == {p for r in ref for p in r["overlaps"]}
    got = {(s, t): w for s, t, w in _rows(cfg, "select source_id,target_id,weight from links")}
    exp = {}
    for t in range(T - 1):
        for w, s, d in link_pair(ref[t]["nodes"], ref[t + 1]["nodes"], 10, 3, time=t):
            exp[(int(s), int(d))] = w
    assert set(got) == set(exp) and len(exp) > 10
    for k in exp:
        assert got[k] == pytest.approx(exp[k], rel=1e-5)


def test_reference_invariants_on_repeated_frames(tmp_path, emu):
    """The reference's own assertions on `timelapse_mock_data` (one frame repeated)."""
    cfg = _config(tmp_path, segmentation__min_area=30, segmentation__max_area=700, segmentation__min_frontier=0.1,
                  linking__max_distance=15, linking__max_neighbors=10)
    fg, ct = synth.make_timelapse((80, 80), 4, seed=9, variant="T", radius=6.0, repeat=True)
    segment(fg, ct, cfg)
    link(cfg)
    nodes = {r[0]: (r[1], loads_node(r[2])) for r in _rows(cfg, "select id, t, pickle from nodes")}
    assert len(nodes) > 20
    # every overlaps row has IoU > 0 and node != ancestor (test_segment_processing.py:59-67)
    pairs = set(_rows(cfg, "select node_id, ancestor_id from overlaps"))
    for a, b in pairs:
        assert a != b
        assert iou(nodes[a][1].bbox, nodes[b][1].bbox, nodes[a][1].mask, nodes[b][1].mask) > 0
    # nodes of one hierarchy do not overlap unless listed (:69-87)
    by_hier = {}
    for i, t, h in _rows(cfg, "select id, t, t_hier_id from nodes"):
        by_hier.setdefault((t, h), []).append(i)
    for ids in by_hier.values():
        for i, a in enumerate(ids):
            for b in ids[i + 1:]:
                if (a, b) in pairs or (b, a) in pairs:
                    continue
                assert iou(nodes[a][1].bbox, nodes[b][1].bbox, nodes[a][1].mask, nodes[b][1].mask) == 0.0
    # linking (test_link_processing.py:40-68)
    links = _rows(cfg, "select source_id, target_id, weight from links")
    by_target = {}
    for s, t, w in links:
        by_target.setdefault(t, []).append(w)
        assert nodes[s][0] + 1 == nodes[t][0]
        d = np.linalg.norm(np.asarray(nodes[s][1].centroid) - np.asarray(nodes[t][1].centroid))
        assert d < 15
    max_t = max(t for t, _ in nodes.values())
    for i, (t, _) in nodes.items():
        if t > 0:
            assert 1 <= len(by_target[i]) <= 10
            assert sum(w == 1.0 for w in by_target[i]) == 1
    sources = {s for s, _, _ in links}
    assert sources == {i for i, (t, _) in nodes.items() if t < max_t}


def test_schema_matches_reference_orm(tmp_path, emu):
    """Column sets of core/database.py:81-124 (asserted like test_segment_processing.py:52-57)."""
    cfg = _config(tmp_path)
    database.create_all(cfg.data_config.database_path)
    with database.connect(cfg.data_config.database_path) as conn:
        cols = lambda t: [(r[1], r[2], r[5]) for r in conn.execute(f"pragma table_info({t})")]  # noqa: E731
        assert cols("nodes") == [
            ("t", "INTEGER", 1), ("id", "BIGINT", 2), ("parent_id", "BIGINT", 0), ("hier_parent_id", "BIGINT", 0),
            ("t_node_id", "INTEGER", 0), ("t_hier_id", "INTEGER", 0), ("z", "FLOAT", 0), ("y", "FLOAT", 0),
            ("x", "FLOAT", 0), ("z_shift", "FLOAT", 0), ("y_shift", "FLOAT", 0), ("x_shift", "FLOAT", 0),
            ("area", "INTEGER", 0), ("frontier", "FLOAT", 0), ("height", "FLOAT", 0), ("selected", "BOOLEAN", 0),
            ("pickle", "BLOB", 0), ("features", "BLOB", 0), ("node_prob", "FLOAT", 0),
            ("segm_annot", "VARCHAR(14)", 0), ("node_annot", "VARCHAR(7)", 0), ("appear_annot", "VARCHAR(7)", 0),
            ("disappear_annot", "VARCHAR(7)", 0), ("division_annot", "VARCHAR(7)", 0)]
        assert [c[0] for c in cols("overlaps")] == ["id", "node_id", "ancestor_id"]
        assert [c[0] for c in cols("links")] == ["id", "source_id", "target_id", "weight", "annotation"]
        tables = {r[0] for r in conn.execute("select name from sqlite_master where type='table'")}
        assert {"nodes", "overlaps", "links", "gt_nodes", "gt_links"} <= tables


def test_errors_and_overwrite(tmp_path, emu):
    cfg = _config(tmp_path)
    fg, ct = synth.make_timelapse((64, 64), 2, seed=1, variant="S", radius=7.0)
    with pytest.raises(ValueError, match="shape must match"):
        segment(fg, ct[:, :32], cfg)
    with pytest.raises(ValueError, match="is empty"):
        link(cfg)
    segment(fg, ct, cfg)
    n0 = _rows(cfg, "select count(*) from nodes")[0][0]
    with pytest.raises(ValueError, match="Duplicated nodes found"):
        segment(fg, ct, cfg)
    segment(fg, ct, cfg, overwrite=True)
    assert _rows(cfg, "select count(*) from nodes")[0][0] == n0
    link(cfg)
    l0 = _rows(cfg, "select count(*) from links")[0][0]
    link(cfg, overwrite=True)
    assert _rows(cfg, "select count(*) from links")[0][0] == l0 > 0
    with pytest.warns(DeprecationWarning):
        segment(detection=fg, edge=ct, config=cfg, overwrite=True)
    with pytest.raises(ValueError, match="exceeds upper bound"):
        segment(fg, ct, cfg, overwrite=True, max_segments_per_time=3)


def test_link_from_database_equals_link_from_device_tables(tmp_path, emu):
    """link() in a fresh process rebuilds the node tables from the pickled blobs."""
    from ultrack_b200 import frames
    cfg = _config(tmp_path)
    fg, ct = synth.make_timelapse((72, 72), 3, seed=4, variant="S", radius=6.0, motion=2.0)
    segment(fg, ct, cfg)
    link(cfg)
    a = sorted(_rows(cfg, "select source_id,target_id,weight from links"))
    frames.STORE.clear()
    link(cfg, overwrite=True)
    b = sorted(_rows(cfg, "select source_id,target_id,weight from links"))
    assert a == b and len(a) > 0


def test_link_uses_shift_columns_and_scale(tmp_path, emu):
    cfg = _config(tmp_path)
    fg, ct = synth.make_timelapse((12, 40, 40), 2, seed=6, variant="S", radius=5.0, motion=1.0)
    segment(fg, ct, cfg)
    rng = np.random.default_rng(0)
    ids = [r[0] for r in _rows(cfg, "select id from nodes where t = 1 order by id")]
    shifts = rng.normal(0, 2.0, size=(len(ids), 3))
    with database.connect(cfg.data_config.database_path) as conn:
        conn.executemany("update nodes set z_shift=?, y_shift=?, x_shift=? where id=?",
                         [(*map(float, s), i) for s, i in zip(shifts, ids)])
    link(cfg, scale=(2.0, 1.0, 1.0))
    prm = SegParams(threshold=0.5, min_area=60, max_area=2000)
    ref = [o_seg(fg[t], ct[t], prm, time=t) for t in range(2)]
    exp = {(int(s), int(d)): w for w, s, d in link_pair(ref[0]["nodes"], ref[1]["nodes"], 10, 3, target_shift=shifts,
                                                        scale=(2.0, 1.0, 1.0))}
    got = {(s, t): w for s, t, w in _rows(cfg, "select source_id,target_id,weight from links")}
    assert set(got) == set(exp) and len(exp) > 5
    for k in exp:
        assert got[k] == pytest.approx(exp[k], rel=1e-5)


def test_empty_frame_warns_and_skips(tmp_path, emu):
    """linking/processing.py:194-209."""
    cfg = _config(tmp_path)
    fg, ct = synth.make_timelapse((64, 64), 3, seed=1, variant="S", radius=7.0)
    fg[1] = 0
    segment(fg, ct, cfg)
    with pytest.warns(RuntimeWarning, match="No segments found"):
        link(cfg)
    assert _rows(cfg, "select count(*) from links")[0][0] == 0


def test_max_noise_replays_numpy_legacy_rng(tmp_path, emu):
    """processing.py:201-209: np.random.seed(t); contours + U(0, max_noise)."""
    cfg = _config(tmp_path, segmentation__max_noise=0.05)
    fg, ct = synth.make_timelapse((64, 64), 2, seed=2, variant="S", radius=7.0)
    segment(fg, ct, cfg)
    prm = SegParams(threshold=0.5, min_area=60, max_area=2000)
    n = 0
    for t in range(2):
        np.random.seed(t)
        noise = np.random.uniform(0, 0.05, size=ct[t].shape).astype(np.result_type(ct.dtype, np.float16))
        ref = o_seg(fg[t], ct[t] + noise, prm, time=t)
        rows = _rows(cfg, f"select id, area, frontier from nodes where t = {t} order by id")
        assert [(r[0], r[1]) for r in rows] == [(m["id"], m["area"]) for m in ref["nodes"]]
        n += len(rows)
    assert n > 0


def test_batch_index_windows(tmp_path, emu):
    """utils/multiprocessing.py:82-105: batch b covers [b*n_workers, (b+1)*n_workers)."""
    cfg = _config(tmp_path, segmentation__n_workers=2, linking__n_workers=2)
    fg, ct = synth.make_timelapse((64, 64), 4, seed=1, variant="S", radius=7.0, motion=1.0)
    segment(fg, ct, cfg, batch_index=0)
    assert {r[0] for r in _rows(cfg, "select distinct t from nodes")} == {0, 1}
    segment(fg, ct, cfg, batch_index=1)
    assert {r[0] for r in _rows(cfg, "select distinct t from nodes")} == {0, 1, 2, 3}
    link(cfg, batch_index=0)
    link(cfg, batch_index=1)
    cfg2 = _config(tmp_path / "b")
    segment(fg, ct, cfg2)
    link(cfg2)
    q = "select source_id,target_id,weight from links order by source_id,target_id"
    assert _rows(cfg, q) == _rows(cfg2, q)


def test_link_color_filter_matches_oracle(tmp_path, emu):
    """link(config, images=...) -- color_filtering_mask, core/linking/processing.py:52-104."""
    from tests import cases
    cfg = _config(tmp_path, linking__z_score_threshold=0.8)
    shape, T = (12, 40, 40), 3
    fg, ct = synth.make_timelapse(shape, T, seed=3, variant="S", radius=5.0, motion=2.0)
    segment(fg, ct, cfg)
    per_t = [cases.color_images(shape, np.float32, seed=20 + t) for t in range(T)]
    images = [np.stack([per_t[t][t % 2][c] for t in range(T)]) for c in range(2)]     # one (T,Z,Y,X) array per channel
    link(cfg, images=images)
    gated = {(s, t): w for s, t, w in _rows(cfg, "select source_id,target_id,weight from links")}
    prm = SegParams(threshold=0.5, min_area=60, max_area=2000)
    ref = [o_seg(fg[t], ct[t], prm, time=t) for t in range(T)]
    exp, plain = {}, 0
    for t in range(T - 1):
        fr = (np.stack([im[t] for im in images], axis=-1), np.stack([im[t + 1] for im in images], axis=-1))
        for w, s, d in link_pair(ref[t]["nodes"], ref[t + 1]["nodes"], 10, 3, time=t, frames=fr, z_score_threshold=0.8):
            exp[(int(s), int(d))] = w
        plain += len(link_pair(ref[t]["nodes"], ref[t + 1]["nodes"], 10, 3, time=t))
    assert set(gated) == set(exp) and 5 < len(exp) < plain
    for k in exp:
        assert gated[k] == pytest.approx(exp[k], rel=1e-5)


def test_match_gt_candidate_links(tmp_path, emu):
    """The GT matcher's compute_spatial_neighbors call (core/match_gt.py:148-168): weight =
    intersection, positive only, rows in gt_nodes / gt_links."""
    import scipy.ndimage as ndi
    from ultrack_b200.match_gt import gt_frame_nodes, match_ground_truth_links
    cfg = _config(tmp_path)
    shape, T = (80, 80), 2
    fg, ct = synth.make_timelapse(shape, T, seed=5, variant="S", radius=7.0, motion=2.0)
    segment(fg, ct, cfg)
    gt = np.stack([ndi.label(np.roll(fg[t] > 0.5, 2, axis=1))[0] for t in range(T)])
    w, s, d = match_ground_truth_links(1, gt, cfg)
    prm = SegParams(threshold=0.5, min_area=60, max_area=2000)
    ref = o_seg(fg[1], ct[1], prm, time=1)
    ndim, labels, bboxes, cens, masks, areas = gt_frame_nodes(gt[1])
    gt_nodes = [dict(id=int(l), centroid=c, bbox=b, mask=m) for l, b, c, m in zip(labels, bboxes, cens, masks)]
    exp = {(int(r[1]), int(r[2])): r[0] for r in link_pair(ref["nodes"], gt_nodes, 10, 3, time=1,
                                                           weight_func="intersection", edge_must_be_positive=True)}
    got = {(int(a), int(b)): float(x) for a, b, x in zip(s, d, w)}
    assert got == exp and len(exp) > 5 and all(v > 0 and v == int(v) for v in exp.values())
    rows = _rows(cfg, "select source_id,target_id,weight,selected from gt_links")
    assert {(a, b): c for a, b, c, _ in rows} == exp
    assert len(_rows(cfg, "select label from gt_nodes where t = 1")) == len(labels)


@pytest.mark.parametrize("shape,channels", [((64, 64), None), ((10, 32, 32), 2)])
def test_segment_properties_features(tmp_path, emu, shape, channels):
    """segment(..., images=, properties=) -- create_feats_callback (processing.py:107-157): float32
    regionprops vectors in `nodes.features`, names in metadata.toml.  skimage is not installable
    here; the expected values are regionprops' definitions restated on the unpickled masks."""
    cfg = _config(tmp_path)
    T = 2
    fg, ct = synth.make_timelapse(shape, T, seed=2, variant="S", radius=6.0 if len(shape) == 2 else 5.0)
    rng = np.random.default_rng(1)
    img = rng.random((T,) + shape + ((channels,) if channels else ())).astype(np.float32)
    props = ["area", "centroid", "intensity_mean", "bbox", "intensity_max", "intensity_min", "extent", "intensity_std"]
    segment(fg, ct, cfg, images=img, properties=props)
    nd = len(shape)
    names = cfg.data_config.metadata["properties"]
    suffix = lambda p: [f"{p}-{c}" for c in range(channels)] if channels else [p]    # noqa: E731
    assert names == (["area"] + [f"centroid-{i}" for i in range(nd)] + suffix("intensity_mean")
                     + [f"bbox-{i}" for i in range(2 * nd)] + suffix("intensity_max") + suffix("intensity_min")
                     + ["extent"] + suffix("intensity_std"))
    rows = _rows(cfg, "select t, pickle, features from nodes order by id")
    assert len(rows) > 5
    for t, blob, feats in rows:
        node = loads_node(blob)
        f = pickle.loads(feats)
        assert f.dtype == np.float32 and f.shape == (len(names),)
        idx = np.nonzero(node.mask)
        coords = np.stack(idx, axis=1) + np.asarray(node.bbox[:nd])
        vals = img[t][tuple(coords.T)].astype(np.float64).reshape(len(coords), -1)
        ext = np.asarray(node.bbox[nd:]) - np.asarray(node.bbox[:nd])
        exp = np.concatenate([[node.area], coords.mean(axis=0), vals.mean(axis=0), node.bbox, vals.max(axis=0),
                              vals.min(axis=0), [node.area / ext.prod()], vals.std(axis=0)]).astype(np.float32)
        assert np.allclose(f, exp, rtol=1e-5, atol=1e-6), (f, exp)
    with pytest.raises(NotImplementedError):
        segment(fg, ct, cfg, properties=["eccentricity"], overwrite=True)


def test_config_mirror_round_trip(tmp_path):
    cfg = MainConfig.model_validate({"segmentation": {"ws_hierarchy": "dynamics", "min_area": 5, "max_area": 10},
                                     "data": {"working_dir": str(tmp_path)}, "tracking": {"appear_weight": -0.5}})
    assert ws_hierarchy_name(cfg.segmentation_config.ws_hierarchy) == "dynamics"
    save_config(cfg, tmp_path / "c.toml")
    again = load_config(tmp_path / "c.toml")
    assert again.model_dump(by_alias=True) == cfg.model_dump(by_alias=True)
    with pytest.raises(ValueError):
        MainConfig.model_validate({"segmentation": {"min_area": 10, "max_area": 5}})
    with pytest.raises(ValueError):
        MainConfig.model_validate({"segmentation": {"ws_hierarchy": "nope"}})
    with pytest.raises(ValueError):
        MainConfig.model_validate({"segmentation": {"unknown_field": 1}})
    assert cfg.data_config.database_path == f"sqlite:///{tmp_path.absolute()}/data.db"


def test_node_blob_format():
    m = np.random.default_rng(0).random((3, 4, 5)) > 0.5
    blob = dumps_node(7, 1000007, 0, int(m.sum()), 0.25, 1.5, np.array([1, 2, 3, 4, 6, 8]), m, np.array([2, 4, 5]))
    ops = [op.name for op, _, _ in __import__("pickletools").genops(blob)]
    assert ops[:6] == ["PROTO", "SHORT_BINUNICODE", "SHORT_BINUNICODE", "STACK_GLOBAL", "EMPTY_TUPLE", "NEWOBJ"]
    assert ops[-2:] == ["BUILD", "STOP"]
    assert b"ultrack.core.segmentation.node" in blob
    node = loads_node(blob)
    assert set(node.__dict__) == {"area", "frontier", "height", "_h_node_index", "_parent", "bbox", "mask", "id",
                                  "time", "centroid"}
    assert np.array_equal(node.mask, m) and node._parent is None
    packed = pack_array(m)
    assert np.array_equal(unpack_array(packed), m)
    # stored Blosc chunk: 16-byte header, nbytes/blocksize/cbytes consistent
    import struct
    ver, verlz, flags, typesize, nbytes, blocksize, cbytes = struct.unpack("<BBBBIII", packed[:16])
    if flags & 0x02:
        assert cbytes == len(packed) == nbytes + 16 and blocksize == nbytes and typesize == 1
        assert np.array_equal(pickle.loads(packed[16:]), m)


def _parse_blosc_chunk(chunk: bytes):
    """Independent reading of the c-blosc2 chunk header (README_CHUNK_FORMAT.rst): byte 0 version,
    1 versionlz, 2 flags (0x1 byte shuffle, 0x2 memcpyed, 0x4 bit shuffle, 0x10 don't split, bits 5-7
    codec), 3 typesize, then little-endian uint32 nbytes, blocksize, cbytes.  flags 0x1|0x4 together
    mark the 32-byte extended (Blosc2) header: bytes 16.. filter codes[6], udcodec, codec meta,
    filter meta[6], reserved, blosc2 flags."""
    import struct
    ver, verlz, flags, typesize, nbytes, blocksize, cbytes = struct.unpack("<BBBBIII", chunk[:16])
    extended = bool(flags & 0x01) and bool(flags & 0x04)
    hdr = 32 if extended else 16
    return dict(version=ver, versionlz=verlz, flags=flags, typesize=typesize, nbytes=nbytes, blocksize=blocksize,
                cbytes=cbytes, header=hdr, memcpyed=bool(flags & 0x02), payload=chunk[hdr:hdr + nbytes])


@pytest.mark.parametrize("shape", [(64, 64), (12, 40, 40)])
def test_bulk_blob_writer_matches_per_node_pickle(emu, shape):
    """`ulk_frame_blobs_to_host` (one kernel for all blobs of a frame) == dumps_node per node, as
    unpickled objects; every mask chunk is a well-formed stored Blosc chunk."""
    from tests.emu_harness import segment_frame
    from ultrack_b200 import _lib
    from ultrack_b200.node import frame_blobs
    fg, ct, _ = synth.make_frame(shape, seed=4, variant="S", radius=5.0, fill=0.4)
    h, nodes = segment_frame(emu, fg, ct, keep_labels=False, min_area=60, max_area=600)
    nd = len(shape)
    buf, off = frame_blobs(emu, h, nodes, 3, nd, 4_000_000)
    assert len(off) == len(nodes["area"]) + 1 > 4
    lo = 3 - nd
    for i in range(len(nodes["area"])):
        blob = bytes(buf[off[i]:off[i + 1]])
        got = loads_node(blob)
        bb = nodes["bbox"][i].astype(np.int64)
        ref = loads_node(dumps_node(int(nodes["hidx"][i]), 4_000_000 + i + 1, 3, int(nodes["area"][i]),
                                    float(nodes["frontier"][i]), float(nodes["height"][i]),
                                    np.concatenate([bb[lo:3], bb[3 + lo:]]), _lib.unpack_mask(nodes, i, nd),
                                    nodes["centroid"][i][lo:].astype(np.int64)))
        assert list(got.__dict__) == list(ref.__dict__)
        for k, b in ref.__dict__.items():
            a = getattr(got, k)
            if isinstance(b, np.ndarray):
                assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b), k
            else:
                assert a == b and type(a) is type(b), (k, a, b)
        # the raw state dict (no __setstate__): mask entry is the Blosc chunk
        class Raw:
            def __setstate__(self, d): self.d = d
        import io
        class U(pickle.Unpickler):
            def find_class(self, module, name):
                return Raw if name == "Node" else super().find_class(module, name)
        chunk = U(io.BytesIO(blob)).load().d["mask"]
        c = _parse_blosc_chunk(chunk)
        assert c["version"] == 2 and c["memcpyed"] and c["typesize"] == 1 and c["header"] == 16
        assert c["cbytes"] == len(chunk) == c["nbytes"] + 16 and c["blocksize"] == c["nbytes"]
        assert np.array_equal(pickle.loads(c["payload"]), got.mask)
    emu.ulk_frame_nodes_free(h)


def test_blosc_chunk_variants_round_trip():
    """Stored chunks in both header forms decode: the 16-byte Blosc1 header this package writes and
    the 32-byte extended header blosc2.pack_array itself emits (flags 0x1|0x4 set)."""
    import struct
    m = np.random.default_rng(2).random((5, 6, 7)) > 0.4
    payload = pickle.dumps(m, pickle.HIGHEST_PROTOCOL)
    n = len(payload)
    c1 = pack_array(m)
    assert _parse_blosc_chunk(c1)["payload"] == payload or np.array_equal(unpack_array(c1), m)
    ext = struct.pack("<BBBBIII", 5, 1, 0x01 | 0x02 | 0x04 | 0x10, 1, n, n, n + 32) + bytes(16) + payload
    c = _parse_blosc_chunk(ext)
    assert c["header"] == 32 and c["memcpyed"] and c["payload"] == payload
    assert np.array_equal(unpack_array(ext), m)
    with pytest.raises(ValueError):
        unpack_array(ext[:20])
