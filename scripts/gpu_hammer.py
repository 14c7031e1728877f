"""Hammer a few parity cases in-process to catch intermittent label mismatches (debugging aid)."""
import os, sys, pathlib, time
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import numpy as np
from tests import cases
from tests.emu_harness import segment_frame
from oracle.segment_oracle import SegParams, segment_frame as o_seg
from ultrack_b200 import _lib

lib = _lib.get_lib()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
todo = []
for kind, hier in (("smooth", "dynamics"), ("quant", "area"), ("rand", "dynamics")):
    for seed in range(2):
        fg, ct = cases.blob_frame((24, 48, 48), seed, kind)
        kw = dict(min_area=40, max_area=1350, ws_hierarchy=hier, min_frontier=0.25 if hier == "dynamics" else 0.0)
        ref = o_seg(fg, ct, SegParams(**kw))
        todo.append((kind, hier, seed, fg, ct, kw, ref))
bad = 0
t0 = time.time()
for rep in range(reps):
    for kind, hier, seed, fg, ct, kw, ref in todo:
        try:
            h, nodes = segment_frame(lib, fg, ct, keep_labels=True, **kw)
        except Exception as e:  # noqa: BLE001
            print("EXC", rep, kind, hier, seed, str(e)[-120:]); bad += 1; sys.exit(1)
        lab = nodes["labels"].reshape(ref["labels"].shape)
        if not np.array_equal(lab, ref["labels"]) or len(nodes["area"]) != len(ref["nodes"]):
            d = lab != ref["labels"]
            print("MISMATCH", rep, kind, hier, seed, "voxels", int(d.sum()), "gpu labels", np.unique(lab[d])[:8], "ref labels",
                  np.unique(ref["labels"][d])[:8], "nregions", lab.max(), ref["labels"].max(), "nodes", len(nodes["area"]), len(ref["nodes"]),
                  "stats", nodes["stats"])
            bad += 1
        lib.ulk_frame_nodes_free(h)
print("env", {k: v for k, v in os.environ.items() if k.startswith("ULK_")}, "reps", reps, "bad", bad, "sec", round(time.time() - t0, 1))
