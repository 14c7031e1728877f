"""GPU exploration: parity spot checks against the oracle, then per-kernel timing on a
larger synthetic frame.  Development aid (not part of the package)."""
import ctypes as C
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from ultrack_b200 import _lib, synth  # noqa: E402
from ultrack_b200.segment import segment_frame_device  # noqa: E402
from tests.emu_harness import compare_nodes, seg_params  # noqa: E402


def parity():
    from oracle.segment_oracle import SegParams, segment_frame as o_seg
    import scipy.ndimage as ndi
    lib = _lib.get_lib()
    nfail = 0
    for seed in range(3):
        for kind in ("rand", "smooth", "quant", "bool"):
            for shape in ((64, 64), (16, 32, 32)):
                rng = np.random.default_rng(seed)
                fgf = ndi.gaussian_filter(rng.random(shape), 2.0)
                fg = fgf > np.quantile(fgf, 0.45)
                if kind == "rand":
                    ct = rng.random(shape).astype(np.float32)
                elif kind == "smooth":
                    ct = ndi.gaussian_filter(rng.random(shape), 1.0).astype(np.float32)
                elif kind == "quant":
                    ct = (np.round(ndi.gaussian_filter(rng.random(shape), 1.0) * 40) / 4 % 1.25).astype(np.float32)
                else:
                    ct = ndi.gaussian_filter(rng.random(shape), 1.0) > 0.52
                for hier in ("area", "dynamics"):
                    kw = dict(min_area=16, max_area=400, ws_hierarchy=hier, min_frontier=0.2 if hier == "dynamics" else 0.0)
                    try:
                        ref = o_seg(fg, ct, SegParams(**kw))
                    except RuntimeError as e:
                        ref = e
                    try:
                        h = segment_frame_device(lib, fg, ct, seg_params(fg_dtype=fg.dtype, **kw), keep_labels=True)
                    except RuntimeError as e:
                        ok = isinstance(ref, RuntimeError)
                        nfail += not ok
                        print("raise", ok, shape, kind, hier, str(e)[:50])
                        continue
                    nodes = _lib.nodes_to_host(lib, h, with_labels=True)
                    lib.ulk_frame_nodes_free(h)
                    try:
                        assert not isinstance(ref, RuntimeError), "oracle raised"
                        assert np.array_equal(nodes["labels"].reshape(ref["labels"].shape), ref["labels"])
                        compare_nodes(nodes, ref, len(shape))
                        print("OK", shape, kind, hier, seed, len(ref["nodes"]))
                    except AssertionError as e:
                        nfail += 1
                        print("FAIL", shape, kind, hier, seed, str(e)[:200])
    print("parity failures:", nfail)
    return nfail


def timing(shape, radius=8.0, reps=3, **kw):
    lib = _lib.get_lib()
    dev = torch.device("cuda:0")
    fg = torch.empty(shape, dtype=torch.float32, device=dev)
    ct = torch.empty(shape, dtype=torch.float32, device=dev)
    shp = (C.c_int64 * 3)(*shape)
    _lib.check(lib, lib.ulk_synth_frame(shp, 0, 7, radius, 0.3, 3.0, 1e-3, C.c_void_p(fg.data_ptr()), C.c_void_p(ct.data_ptr()), None))
    prm = seg_params(fg_dtype=np.float32, **kw)
    torch.cuda.synchronize()
    V = shape[0] * shape[1] * shape[2]
    for r in range(reps):
        if r == reps - 1:
            lib.ulk_profile_reset()
            lib.ulk_profile_enable(1)
        t0 = time.perf_counter()
        h = segment_frame_device(lib, fg, ct, prm)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        f = h.contents
        print(f"shape {shape} rep {r}: {dt*1e3:.1f} ms  {V/dt/1e6:.1f} Mvox/s nodes {f.n_nodes} hier {f.n_hier} "
              f"basins {f.stats.n_basins} rag {f.stats.n_rag_edges} flow_rounds {f.stats.flow_rounds} k3 {f.stats.k3_rounds} "
              f"plateau {f.stats.plateau_voxels} maskwords {f.mask_words}")
        lib.ulk_frame_nodes_free(h)
    lib.ulk_profile_enable(0)
    rep = _lib.profile_report(lib)
    tot = sum(v[1] for v in rep.values())
    print(f"kernel time total {tot:.1f} ms over {sum(v[0] for v in rep.values())} launches")
    for k, v in sorted(rep.items(), key=lambda kv: -kv[1][1])[:25]:
        print(f"  {v[1]:9.3f} ms {v[0]:5d}x  {k}")
    free, total = torch.cuda.mem_get_info()
    print("mem used GB", (total - free) / 1e9)
    return rep


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    nf = parity()
    timing((1, 2048, 2048), radius=12.0, min_area=100, max_area=100000)
    timing((128, 512, 512), min_area=100, max_area=10000)
    if len(sys.argv) > 1 and sys.argv[1] == "big":
        timing((256, 1024, 1024), min_area=100, max_area=10000, reps=2)
    sys.exit(1 if nf else 0)
