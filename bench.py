#!/usr/bin/env python
"""bench.py -- voxels/s of segment+link (and links/s) on N B200s of one node.

A step = one pass of the hot path over one frame: `ulk_segment_frame` on a synthetic frame
plus `ulk_link_frames` of the previous frame's node table against the new one.  Frames are
sharded by timepoint across ranks with no data-path collective (weak scaling: every rank
runs the same number of steps on its own frames); the shard-boundary exchange of link() is a
per-shard constant that bench_boundary() reports separately.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c4|c3|c2] [--impl reference]

`--impl reference` times the CPU restatement of the reference's path (oracle/, the reference
itself is not installable here: it needs higra) on the box's host cores on a bounded sample.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    # name: (frame shape, cell radius, BASELINE.json config it is the per-GPU unit of)
    "c5": ((448, 1536, 1536), 8.0, "configs[4] Zebrahub-scale 3D 448x1536x1536 frame, max_distance=15"),
    "c4": ((256, 1024, 1024), 8.0, "configs[3]-sized 3D 256x1024x1024 frame"),
    "c3": ((128, 512, 512), 8.0, "configs[2] 3D 128x512x512 frame, min_area=100 max_area=10000"),
    "c2": ((1, 2048, 2048), 12.0, "configs[1] 2D 2048x2048 frame, max_neighbors=5"),
}
SEG = dict(threshold=0.5, min_area=100, max_area=10000, min_area_factor=4.0, min_frontier=0.0, ws_hierarchy="area")
LINK = dict(max_distance=15.0, max_neighbors=5, distance_weight=0.0)
SYNTH = dict(seed=7, fill=0.3, motion=3.0, noise=1e-3)
METRIC = "voxels_per_sec_segment_link"


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None
        self.nvml = None

    def _nvml_loop(self, pynvml, h):
        R = {"hw_slowdown": pynvml.nvmlClocksEventReasonHwSlowdown,
             "hw_thermal_slowdown": pynvml.nvmlClocksEventReasonHwThermalSlowdown,
             "sw_thermal_slowdown": pynvml.nvmlClocksEventReasonSwThermalSlowdown,
             "sw_power_cap": pynvml.nvmlClocksEventReasonSwPowerCap}
        mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
        while not self._stop.is_set():
            sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            bits = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
            self.rows.append([str(sm), str(mx)] + ["Active" if bits & R[n] else "Not Active" for n in
                                                   ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")])
            self._stop.wait(0.05)

    def start(self):
        if os.environ.get("ULK_BENCH_NO_CLOCKS"):      # diagnosis only: no sampling at all
            return
        # in-process NVML (two light queries per sample); a spawned `nvidia-smi -lms` re-initialises
        # the driver query path every 100 ms and was seen to stall kernel launches for > 100 ms
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            self._stop = threading.Event()
            self.nvml = threading.Thread(target=self._nvml_loop, args=(pynvml, h), daemon=True)
            self.nvml.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index),
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml is not None:
            self._stop.set()
            self.nvml.join(1.0)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            time.sleep(0.15)
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU arm
REF_NOTEBOOK_MVOX_PER_WORKER = 6.52    # ultrack's own figure: examples/zebrahub/zebrahub.ipynb:407


def _cpu_worker(conn, slot: int, sshape, radius: float, nframes: int):
    """One host core of the CPU arm.  Everything that is not the path is done BEFORE the clock:
    the worker generates its `nframes` sample frames and segments frame 0 (so that step k can link
    frame k+1 to its predecessor), then reports ready.  A step = segment ONE frame + link ONE pair."""
    import time as _t

    from oracle.link_oracle import link_pair
    from oracle.segment_oracle import SegParams, segment_frame
    from ultrack_b200 import synth
    prm = SegParams(threshold=SEG["threshold"], min_area=SEG["min_area"], max_area=SEG["max_area"],
                    min_area_factor=SEG["min_area_factor"], min_frontier=SEG["min_frontier"],
                    ws_hierarchy=SEG["ws_hierarchy"])
    frames = []
    for j in range(nframes):
        fg, ct, _ = synth.make_frame(sshape, t=slot * nframes + j, seed=SYNTH["seed"], variant="S", radius=radius,
                                     fill=SYNTH["fill"], motion=SYNTH["motion"], noise=SYNTH["noise"])
        frames.append((fg, ct))
    prev = segment_frame(frames[0][0], frames[0][1], prm, time=0)["nodes"]
    conn.send("ready")
    while True:
        k = conn.recv()
        if k is None:
            break
        fg, ct = frames[1 + k % (nframes - 1)]
        t0 = _t.perf_counter()
        cur = segment_frame(fg, ct, prm, time=k + 1)["nodes"]
        links = link_pair(prev, cur, LINK["max_distance"], LINK["max_neighbors"], LINK["distance_weight"], time=k)
        dt = _t.perf_counter() - t0
        prev = cur
        conn.send((dt, 0 if links is None else len(links)))
    conn.close()


def _ws_stats(lib):
    out = (C.c_int64 * 3)()
    lib.ulk_workspace_stats(out)
    return [int(v) for v in out]


def cpu_sample_shape(shape):
    """Bounded sample of the workload for the CPU arm: a sub-volume of the same synthetic recipe."""
    if shape[0] == 1:
        return (384, 384)
    return (40, 128, 128)


def run_cpu_arm(shape, radius, steps: int, warmup: int, max_workers: int = 0):
    """The restated reference path on the host cores, one persistent worker process per core over
    its own timepoints (the reference's own parallelism: multiprocessing over timepoints,
    utils/multiprocessing.py:26-56).  The wall clock covers segment + link only: inputs are
    generated and the first frame is segmented before it starts."""
    import multiprocessing as mp
    from oracle.build import build as build_oracle
    build_oracle()
    cores = max_workers or min(os.cpu_count() or 1, 64)
    sshape = cpu_sample_shape(shape)
    vox = 1
    for s in sshape:
        vox *= s
    steps, warmup = max(steps, 1), max(warmup, 0)
    nframes = 1 + min(steps + warmup, 4)           # distinct frames per worker (cycled when more steps are asked)
    ctx = mp.get_context("spawn")
    workers = []
    for slot in range(cores):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_cpu_worker, args=(b, slot, sshape, radius, nframes), daemon=True)
        p.start()
        workers.append((p, a))
    for _, c in workers:
        assert c.recv() == "ready"

    def one(k):
        for _, c in workers:
            c.send(k)
        return [c.recv() for _, c in workers]

    for k in range(warmup):
        one(k)
    t0 = time.perf_counter()
    n_links, busy = 0, 0.0
    for k in range(steps):
        res = one(warmup + k)
        n_links += sum(r[1] for r in res)
        busy += sum(r[0] for r in res)
    wall = time.perf_counter() - t0
    for p, c in workers:
        c.send(None)
    for p, _ in workers:
        p.join(5)
    vps = vox * cores * steps / wall
    return {"voxels_per_sec": vps, "links_per_sec": n_links / wall, "cores": cores, "sample_shape": list(sshape),
            "per_core_voxels_per_sec": vps / cores, "worker_busy_fraction": busy / (wall * cores),
            "reference_notebook_voxels_per_sec_per_worker": REF_NOTEBOOK_MVOX_PER_WORKER * 1e6,
            "sample": f"{steps} timed step(s) after {warmup} warm-up x {cores} worker process(es), each step = segment ONE "
                      f"{'x'.join(map(str, sshape))} sub-volume of the workload's synthetic recipe + link it to its "
                      f"predecessor; frames pre-generated and frame 0 segmented before the clock; wall {wall:.1f}s; "
                      f"port runs at {vps / cores / 1e6:.2f} Mvoxel/s/core, the reference's own notebook shows "
                      f"{REF_NOTEBOOK_MVOX_PER_WORKER} Mvoxel/s/worker (examples/zebrahub/zebrahub.ipynb:407)",
            "ms_per_step": wall * 1e3 / steps}


def make_config(name: str):
    """`config` of the JSON line: the same dict on the GPU arm and the reference arm."""
    shape, radius, desc = WORKLOADS[name]
    return {"workload": desc, "frame_shape": list(shape), "segmentation": SEG, "linking": LINK,
            "synthetic": dict(SYNTH, radius=radius), "l2": "inputs larger than L2 (no flush needed)",
            "parallelism": "timepoint shards, one per GPU (reference arm: one per host core)"}


# ----------------------------------------------------------------------------- extra legs of the GPU arm
def run_api_leg(lib, torch, dist, rank, world, local_rank, frames_per_rank: int = 4):
    """`e2e_api`: the PUBLIC API end to end -- ultrack_b200.segment() + link() on a C3-size
    timelapse held in pinned host memory, rows + pickle blobs into a sqlite file on local disk
    (what the reference's own timings include: SQL and blobs).  Timepoints are sharded over the
    ranks exactly as a user's torchrun job would; wall clock, max over ranks."""
    import shutil
    import tempfile

    import numpy as np

    import ultrack_b200
    from ultrack_b200 import frames as frames_mod
    seg_mod = sys.modules["ultrack_b200.segment"]      # the module (the package re-exports the function under this name)
    shape, radius, _ = WORKLOADS["c3"]
    T = frames_per_rank * world
    dev = torch.device("cuda", local_rank)
    fg_h = torch.empty((T,) + shape, dtype=torch.float32).pin_memory()
    ct_h = torch.empty((T,) + shape, dtype=torch.float32).pin_memory()
    shp = (C.c_int64 * 3)(*shape)
    fg_d = torch.empty(shape, dtype=torch.float32, device=dev)
    ct_d = torch.empty(shape, dtype=torch.float32, device=dev)
    for t in range(T):
        _lib_check(lib, lib.ulk_synth_frame(shp, t, SYNTH["seed"], radius, SYNTH["fill"], SYNTH["motion"], SYNTH["noise"],
                                            C.c_void_p(fg_d.data_ptr()), C.c_void_p(ct_d.data_ptr()), None))
        torch.cuda.synchronize()
        fg_h[t].copy_(fg_d)
        ct_h[t].copy_(ct_d)
    del fg_d, ct_d
    wd = [None]
    if rank == 0:
        wd[0] = tempfile.mkdtemp(prefix="ulk_bench_api_", dir="/tmp")
    if world > 1:
        dist.broadcast_object_list(wd, src=0)
    cfg = ultrack_b200.MainConfig.model_validate({
        "data": {"working_dir": wd[0]},
        "segmentation": {k: SEG[k] for k in ("threshold", "min_area", "max_area", "min_area_factor", "min_frontier", "ws_hierarchy")},
        "linking": dict(LINK)})
    fg_np, ct_np = fg_h.numpy(), ct_h.numpy()

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # warm-up on a throw-away database (pool growth, sqlite file creation)
    ultrack_b200.segment(fg_np[:world], ct_np[:world], cfg, overwrite=True)
    for k in seg_mod.TIMERS:
        seg_mod.TIMERS[k] = 0.0
    sync()
    t0 = time.perf_counter()
    ultrack_b200.segment(fg_np, ct_np, cfg, overwrite=True)
    sync()
    t1 = time.perf_counter()
    ultrack_b200.link(cfg)
    sync()
    t2 = time.perf_counter()
    tt = torch.tensor([t1 - t0, t2 - t1], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    out = None
    if rank == 0:
        import sqlite3
        db = os.path.join(wd[0], "data.db")
        conn = sqlite3.connect(db)
        n_nodes = conn.execute("select count(*) from nodes").fetchone()[0]
        n_links = conn.execute("select count(*) from links").fetchone()[0]
        blob_bytes = conn.execute("select sum(length(pickle)) from nodes").fetchone()[0]
        conn.close()
        seg_s, link_s = float(tt[0]), float(tt[1])
        V = shape[0] * shape[1] * shape[2]
        split = {k: round(v * 1e3 / frames_per_rank, 2) for k, v in seg_mod.TIMERS.items()}
        out = {"value": V * T / (seg_s + link_s), "unit": "voxels/s", "frames": T, "frame_shape": list(shape),
               "segment_s": seg_s, "link_s": link_s, "nodes": n_nodes, "links": n_links, "blob_bytes": blob_bytes,
               "db_bytes": os.path.getsize(db),
               "rank0_ms_per_frame": dict(split, note="segment() wall split on rank 0: kernels incl. upload wait | node table D2H | "
                                                      "blob kernel + row assembly | sqlite insert (writer thread, overlapped with the next frame)"),
               "api": "ultrack_b200.segment(fg, ct, cfg, overwrite=True) + ultrack_b200.link(cfg): pinned host float32 "
                      "timelapse in, nodes / overlaps / links rows with pickle blobs (stored Blosc chunks: blosc2 is not "
                      "installable here, masks are 1 byte per bbox voxel) into sqlite on local disk"}
    frames_mod.STORE.clear()
    sync()
    if rank == 0:
        shutil.rmtree(wd[0], ignore_errors=True)
    return out


def _lib_check(lib, rc):
    from ultrack_b200 import _lib
    _lib.check(lib, rc)


def run_h2d_ceiling(torch, dist, world, local_rank, host_bufs, reps: int = 3):
    """Plain pinned host->device copies of one step's inputs, all ranks at once, no kernels: the
    box's ceiling for any end-to-end number that starts from host memory."""
    dev = torch.device("cuda", local_rank)
    dst = [torch.empty_like(b, device=dev) for b in host_bufs]
    nbytes = sum(b.numel() * b.element_size() for b in host_bufs)
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for d, b in zip(dst, host_bufs):
            d.copy_(b, non_blocking=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(reps):
            for d, b in zip(dst, host_bufs):
                d.copy_(b, non_blocking=True)
        e1.record(st)
        torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms[0]) / reps
    return {"bytes_per_step": nbytes, "ms_per_step": ms, "gbps_per_rank": nbytes / ms / 1e6, "gbps_total": nbytes * world / ms / 1e6}


def run_boundary_leg(lib, torch, dist, rank, world, handle, ids_len, ndim, lprm, sp):
    """The one exchange of the sharded path: rank g+1 sends the node table of the first frame of
    its block to rank g (frames.send_frame / recv_frame: NCCL send/recv of device buffers over
    NVLink), and rank g links its last frame against it.  Timed on the device, max over ranks."""
    import numpy as np

    from ultrack_b200 import _lib, frames as frames_mod
    dev = torch.device("cuda", torch.cuda.current_device())
    mine = frames_mod.DeviceFrame(lib, handle, np.arange(ids_len, dtype=np.int64), ndim)
    f = handle.contents
    nbytes = int(f.n_nodes) * (10 * 4 + 2 * 8) + int(f.mask_words) * 4
    res = {}
    for rep in range(2):           # first round warms NCCL's p2p channels up
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        got = None
        # even ranks receive first, odd ranks send first: every send has a matching receive posted
        for phase in range(2):
            if (rank % 2 == 0) == (phase == 0):
                if rank + 1 < world:
                    got = frames_mod.recv_frame(lib, rank + 1)
            else:
                if rank > 0:
                    frames_mod.send_frame(mine, rank - 1)
        e1.record()
        n_links = 0
        if got is not None:
            out = C.POINTER(_lib.Links)()
            _lib.check(lib, lib.ulk_link_frames(mine.handle, got.handle, None, C.byref(lprm), None, C.byref(out)))
            n_links = int(out.contents.n_links)
            lib.ulk_links_free(out)
            got.free()
        e2.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1), e1.elapsed_time(e2), float(n_links)], dtype=torch.float64, device=dev)
        mx = t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        res = {"nranks": world, "messages": world - 1, "bytes_per_message": nbytes, "exchange_ms": float(mx[0]),
               "boundary_link_ms": float(mx[1]), "boundary_links": int(mx[2]),
               "how": "frames.send_frame/recv_frame (torch.distributed NCCL send/recv of device buffers) + ulk_link_frames "
                      "of the received table; CUDA events, max over ranks; second of two rounds"}
    mine.handle = None              # the caller still owns the handle
    return res


# ----------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("ULK_BENCH_WORKLOAD", "auto"))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-api", action="store_true", help="skip the segment()+link()+sqlite leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: everything libraries print on fd 1 (NCCL's version
    # banner when a communicator comes up, ...) goes to stderr, the line is written to the real fd
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line: dict) -> None:
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    if args.impl == "reference":
        if rank != 0:
            return
        name = "c5" if args.workload == "auto" else args.workload
        shape, radius, desc = WORKLOADS[name]
        res = run_cpu_arm(shape, radius, max(args.steps, 1), args.warmup)
        cb = {"value": res["voxels_per_sec"], "unit": "voxels/s", "cores": res["cores"], "kind": "port",
              "sample": res["sample"], "sample_shape": res["sample_shape"],
              "per_core_voxels_per_sec": res["per_core_voxels_per_sec"],
              "worker_busy_fraction": res["worker_busy_fraction"],
              "reference_notebook_voxels_per_sec_per_worker": res["reference_notebook_voxels_per_sec_per_worker"]}
        line = {
            "impl": "reference", "metric": METRIC, "value": res["voxels_per_sec"], "unit": "voxels/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(make_config(name), sample_shape=res["sample_shape"]),
            "links_per_sec": res["links_per_sec"],
            "cpu_baseline": cb,
            "e2e": {"value": res["voxels_per_sec"], "unit": "voxels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        emit(line)
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    from ultrack_b200 import _lib, bytes_model
    from ultrack_b200.capi import seg_params

    torch.cuda.set_device(local_rank)
    # the rank's host threads (and so its pinned staging buffers) stay on the NUMA node next to
    # its GPU: the end-to-end figure is bound by host->device copies
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(vis.split(",")[local_rank]) if vis and vis.split(",")[local_rank].isdigit() else local_rank
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(phys))
    except Exception:  # noqa: BLE001
        pass
    if world > 1:
        # NCCL's version / debug lines go to stderr: stdout carries exactly one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = _lib.get_lib()
    _lib.check(lib, lib.ulk_set_device(local_rank))
    dev = torch.device("cuda", local_rank)

    name = args.workload
    if name == "auto":
        free, _ = torch.cuda.mem_get_info()
        name = "c5" if free > 150e9 else "c4"
    shape, radius, desc = WORKLOADS[name]
    V = shape[0] * shape[1] * shape[2]
    shp = (C.c_int64 * 3)(*shape)
    stream = torch.cuda.Stream()
    sp = C.c_void_p(stream.cuda_stream)

    # two device-resident synthetic frames per rank (cells move between them); every rank owns
    # its own timepoints (t = 2*rank, 2*rank+1): the inputs are far larger than L2 (126 MB)
    frames_in = []
    for j in range(2):
        fg = torch.empty(shape, dtype=torch.float32, device=dev)
        ct = torch.empty(shape, dtype=torch.float32, device=dev)
        _lib.check(lib, lib.ulk_synth_frame(shp, 2 * rank + j, SYNTH["seed"], radius, SYNTH["fill"], SYNTH["motion"],
                                            SYNTH["noise"], C.c_void_p(fg.data_ptr()), C.c_void_p(ct.data_ptr()), sp))
        frames_in.append((fg, ct))
    prm = seg_params(fg_dtype=np.float32, **SEG)
    lprm = _lib.LinkParams(LINK["max_distance"], LINK["distance_weight"], LINK["max_neighbors"], 0,
                           (C.c_double * 3)(1, 1, 1))

    def seg(j, host=None):
        h = C.POINTER(_lib.FrameNodes)()
        if host is None:
            fg, ct = frames_in[j]
            rc = lib.ulk_segment_frame(C.c_void_p(fg.data_ptr()), 5, C.c_void_p(ct.data_ptr()), 5, shp, C.byref(prm),
                                       0, sp, C.byref(h))
        else:
            fg, ct = host[j]
            rc = lib.ulk_segment_frame_host(C.c_void_p(fg.data_ptr()), 5, C.c_void_p(ct.data_ptr()), 5, shp,
                                            C.byref(prm), 0, sp, C.byref(h))
        _lib.check(lib, rc)
        return h

    def lnk(a, b):
        out = C.POINTER(_lib.Links)()
        _lib.check(lib, lib.ulk_link_frames(a, b, None, C.byref(lprm), sp, C.byref(out)))
        return out

    def step(i, prev, host=None, fetch=False):
        cur = seg(i % 2, host)
        links = lnk(prev, cur)
        n_links, n_pairs = int(links.contents.n_links), int(links.contents.n_pairs)
        d2h = 0
        if fetch:   # the result a user reads back: node table + links
            nodes = _lib.nodes_to_host(lib, cur, stream.cuda_stream)
            s = np.zeros(n_links, np.int32); t = np.zeros(n_links, np.int32); w = np.zeros(n_links, np.float64)
            p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
            _lib.check(lib, lib.ulk_links_to_host(links, p(s), p(t), p(w), sp))
            d2h = sum(v.nbytes for v in nodes.values() if isinstance(v, np.ndarray)) + s.nbytes + t.nbytes + w.nbytes
        lib.ulk_links_free(links)
        lib.ulk_frame_nodes_free(prev)
        return cur, n_links, n_pairs, d2h

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        prev = seg(1)
        # per-kernel events are on during the warm-up too, so the timed steps take exactly the
        # same host path (event pool already created); the reset drops the warm-up's records
        lib.ulk_profile_enable(1)
        for i in range(args.warmup):
            prev, *_ = step(i, prev)
        # ---- timed region: K steps, device-resident inputs, per-kernel events on
        lib.ulk_profile_reset()
        launches0 = lib.ulk_launch_count()
        clocks = ClockSampler(local_rank)
        clocks.start()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        tot_links = tot_pairs = 0
        step_ev = []
        for i in range(args.steps):
            prev, nl, npairs, _ = step(i, prev)
            tot_links += nl
            tot_pairs += npairs
            step_ev.append(torch.cuda.Event(enable_timing=True))
            step_ev[-1].record(stream)
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        ms_steps = [round(a.elapsed_time(b), 3) for a, b in zip([e0] + step_ev[:-1], step_ev)]
        clk = clocks.stop()
        lib.ulk_profile_enable(0)
        launches = lib.ulk_launch_count() - launches0
        prof = _lib.profile_report(lib)
        f = prev.contents
        stats = dict(n_nodes=int(f.n_nodes), n_overlaps=int(f.n_overlaps), mask_words=int(f.mask_words),
                     n_hier=int(f.n_hier), n_basins=int(f.stats.n_basins), n_rag_edges=int(f.stats.n_rag_edges))
        n_fg = int(f.stats.n_foreground)
        n_fg_a = int(f.stats.pass_a_voxels)
        stats["n_foreground"], stats["pass_a_voxels"] = n_fg, n_fg_a
        stats["k3_voxels"] = int(f.stats.k3_voxels)
        stats["flow_rounds"], stats["k3_rounds"] = int(f.stats.flow_rounds), int(f.stats.k3_rounds)
        stats["plateau_voxels"] = int(f.stats.plateau_voxels)

        # ---- e2e: the same steps from pinned HOST buffers through the ingest API segment() uses
        # (ulk_upload_begin of frame i+1 before ulk_segment_frame_uploaded of frame i), node
        # table + links read back to the host every step
        e2e = None
        e2e_narrow = None
        h2d_ceiling = {}
        if not args.no_e2e:
            def run_e2e(host, codes, prm_, prev):
                def begin(j):
                    up = C.c_void_p()
                    fg, ct = host[j % 2]
                    _lib.check(lib, lib.ulk_upload_begin(C.c_void_p(fg.data_ptr()), codes[0], C.c_void_p(ct.data_ptr()),
                                                         codes[1], shp, C.byref(up)))
                    return up

                def step_up(up, prev):
                    h = C.POINTER(_lib.FrameNodes)()
                    _lib.check(lib, lib.ulk_segment_frame_uploaded(up, C.byref(prm_), 0, sp, C.byref(h)))
                    links = lnk(prev, h)
                    n_links = int(links.contents.n_links)
                    nodes = _lib.nodes_to_host(lib, h, stream.cuda_stream, pinned=True)
                    s = np.zeros(n_links, np.int32); t = np.zeros(n_links, np.int32); w = np.zeros(n_links, np.float64)
                    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
                    _lib.check(lib, lib.ulk_links_to_host(links, p(s), p(t), p(w), sp))
                    d2h = sum(v.nbytes for v in nodes.values() if isinstance(v, np.ndarray)) + s.nbytes + t.nbytes + w.nbytes
                    lib.ulk_links_free(links)
                    lib.ulk_frame_nodes_free(prev)
                    return h, d2h

                ke = max(1, min(args.steps, 5))
                up = begin(0)
                nxt = begin(1)
                prev, _ = step_up(up, prev)          # untimed: fills the pipeline
                up = nxt
                barrier()
                e0.record(stream)
                d2h_tot = 0
                for i in range(ke):
                    nxt = begin(i)                   # frame i+2 of the sequence starts copying now
                    prev, d2h = step_up(up, prev)
                    up = nxt
                    d2h_tot += d2h
                e1.record(stream)
                barrier()
                lib.ulk_upload_free(up)
                return prev, (ke, e0.elapsed_time(e1), d2h_tot // ke)

            # (a) the inputs as BASELINE.json's config states them: float32 + float32
            host = [(fg.cpu().pin_memory(), ct.cpu().pin_memory()) for fg, ct in frames_in]
            prev, e2e = run_e2e(host, (5, 5), prm, prev)
            h2d_ceiling["f32+f32"] = run_h2d_ceiling(torch, dist, world, local_rank, list(host[0]))
            del host
            # (b) the dtypes the reference's own Zebrahub example feeds: bool
            # foreground + float16 contours, 3 B/voxel instead of 8
            host = [((fg > SEG["threshold"]).to(torch.uint8).cpu().pin_memory(), ct.to(torch.float16).cpu().pin_memory())
                    for fg, ct in frames_in]
            prm_u8 = seg_params(fg_dtype=np.uint8, **SEG)
            prev, e2e_narrow = run_e2e(host, (0, 4), prm_u8, prev)
            h2d_ceiling["u8+f16"] = run_h2d_ceiling(torch, dist, world, local_rank, list(host[0]))
            del host
    boundary = None
    if world > 1:
        boundary = run_boundary_leg(lib, torch, dist, rank, world, prev, int(prev.contents.n_nodes), 3, lprm, sp)
    lib.ulk_frame_nodes_free(prev)
    api = None
    if not args.no_api:
        del frames_in
        torch.cuda.empty_cache()
        api = run_api_leg(lib, torch, dist, rank, world, local_rank)

    t = torch.tensor([ms, e2e[1] if e2e else 0.0, e2e_narrow[1] if e2e_narrow else 0.0], dtype=torch.float64, device=dev)
    cnt = torch.tensor([tot_links], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_max, ems_max, ems2_max = float(t[0]), float(t[1]), float(t[2])
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if pk.exists() else "B200_PROFILING.md fallback 6650 GB/s (of fallback)"

    value = V * args.steps * world / (ms_max * 1e-3)
    kern_ms = sum(v[1] for v in prof.values())
    top = sorted(prof.items(), key=lambda kv: -kv[1][1])
    # the dominant kernel = the one with the largest share of the timed region that is one of
    # this library's voxel kernels (it has an entry in the byte model)
    dom_name, (dom_cnt, dom_ms) = next(((k, v) for k, v in top if bytes_model.kernel_bytes(k, V, n_fg, n_fg_a)), top[0])
    dom_bytes = bytes_model.kernel_bytes(dom_name, V, n_fg, n_fg_a)
    dom_avg_ms = dom_ms / dom_cnt
    traffic = None
    tf = ROOT / "profiles" / "ncu_traffic.json"     # dram bytes per launch from the committed ncu launch list
    dram_step = None
    if tf.exists():
        tj = json.loads(tf.read_text()).get(name, {})
        traffic = tj.get(dom_name.split("<")[0])
        # measured DRAM bytes of one step = sum over kernels of (ncu bytes per launch x launches per step)
        short = lambda k: k.split("<")[0].split("::")[-1]   # noqa: E731
        seen = [(k, v) for k, v in prof.items() if short(k) in tj or k in tj]
        if seen:
            dram_step = sum((tj.get(k) or tj.get(short(k))) * v[0] for k, v in seen) / max(args.steps, 1)
    roofline = {
        "bound": "hbm", "kernel": dom_name, "launches": dom_cnt, "avg_ms": dom_avg_ms,
        "share_of_kernel_time": dom_ms / kern_ms if kern_ms else None,
        "achieved": (dom_bytes / (dom_avg_ms * 1e-3) / 1e9) if dom_bytes else None, "peak": peak_gbs, "unit": "GB/s",
        "frac": (dom_bytes / (dom_avg_ms * 1e-3) / 1e9 / peak_gbs) if dom_bytes else None,
        "algorithmic_bytes_per_launch": dom_bytes, "traffic": traffic, "peak_source": peak_src,
        "n_foreground": n_fg, "pass_a_voxels": n_fg_a,
    }
    b_frame = bytes_model.frame_bytes(V, 4, 4, n_fg, stats["n_nodes"], stats["mask_words"], stats["n_overlaps"])
    b_link = bytes_model.link_bytes(tot_links // max(args.steps, 1), tot_pairs // max(args.steps, 1),
                                    stats["mask_words"] / max(stats["n_nodes"], 1))
    pipe_gbs = (b_frame + b_link) * args.steps / (ms * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": value, "unit": "voxels/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "ms_steps_rank0": ms_steps,
        "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(name), "frame_stats": stats,
        "links_per_sec": float(cnt[0]) / (ms_max * 1e-3),
        "gpu_launches": int(launches),
        "workspace_bytes": dict(zip(("slab", "peak_in_use", "outside_slab"), _ws_stats(lib))),
        "clocks": clk,
        "pipeline_roofline": {"algorithmic_bytes_per_step": b_frame + b_link, "achieved": pipe_gbs, "peak": peak_gbs,
                              "unit": "GB/s", "frac": pipe_gbs / peak_gbs,
                              "dram_bytes_per_step_ncu": dram_step,
                              "dram_over_algorithmic": (dram_step / (b_frame + b_link)) if dram_step else None,
                              "kernel_time_ms_per_step": kern_ms / args.steps,
                              "top_kernels_ms_per_step": {k: round(v[1] / args.steps, 3) for k, v in top[:80]}},
        "roofline": roofline,
    }
    if e2e:
        ke = e2e[0]
        line["e2e"] = {"value": V * ke * world / (ems_max * 1e-3), "unit": "voxels/s", "steps": ke,
                       "ms_per_step": ems_max / ke, "h2d_bytes_per_step": 2 * 4 * V, "d2h_bytes_per_step": e2e[2],
                       "api": "ulk_upload_begin (next frame, copy stream) + ulk_segment_frame_uploaded + ulk_frame_nodes_to_host + ulk_link_frames + ulk_links_to_host, pinned host buffers; one upload per step inside the timed region"}
    if e2e_narrow:
        ke = e2e_narrow[0]
        line["e2e_u8f16"] = {"value": V * ke * world / (ems2_max * 1e-3), "unit": "voxels/s", "steps": ke,
                             "ms_per_step": ems2_max / ke, "h2d_bytes_per_step": 3 * V, "d2h_bytes_per_step": e2e_narrow[2],
                             "inputs": "uint8 foreground + float16 contours (the reference's Zebrahub example dtypes), same API as e2e"}
    if h2d_ceiling:
        for key, e in (("f32+f32", line.get("e2e")), ("u8+f16", line.get("e2e_u8f16"))):
            if e and key in h2d_ceiling:
                h2d_ceiling[key]["e2e_fraction_of_ceiling"] = h2d_ceiling[key]["ms_per_step"] / e["ms_per_step"]
        line["h2d_ceiling"] = dict(h2d_ceiling, how="pinned host->device cudaMemcpyAsync of one step's inputs, all ranks "
                                                    "concurrently, no kernels; CUDA events, max over ranks")
    if boundary:
        line["boundary"] = boundary
    if api:
        line["e2e_api"] = api
    if not args.no_cpu_baseline and world == 1:
        res = run_cpu_arm(shape, radius, 3, 1)
        line["cpu_baseline"] = {"value": res["voxels_per_sec"], "unit": "voxels/s", "cores": res["cores"], "kind": "port",
                                "sample": res["sample"], "sample_shape": res["sample_shape"],
                                "links_per_sec": res["links_per_sec"],
                                "per_core_voxels_per_sec": res["per_core_voxels_per_sec"],
                                "reference_notebook_voxels_per_sec_per_worker": res["reference_notebook_voxels_per_sec_per_worker"]}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
