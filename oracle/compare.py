"""id-keyed comparison of a device node table with the oracle's output -- TEST INFRASTRUCTURE ONLY
(SURVEY.md section 8a parity definition); used by tests/ and __graft_entry__.smoke()."""
from __future__ import annotations

from typing import Dict

import numpy as np

from ultrack_b200 import _lib


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
