#!/usr/bin/env python
"""Summarises the ncu outputs of scripts/gpu_evidence.sh into profiles/.

  launches_<wl>.csv  (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum)
      -> profiles/<tag>_ncu_launches_<wl>.csv   per kernel: launches, total ms, share, DRAM bytes per launch
      -> profiles/ncu_traffic.json              {workload: {kernel: dram bytes per launch}}  (read by bench.py)
  full_<wl>.ncu-rep  (ncu --set full) -> profiles/<tag>_ncu_full_<wl>.csv   selected raw metrics per captured launch

usage: python scripts/ncu_summarise.py <tag> [gpurun_out]
"""
import collections
import csv
import json
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
tag = sys.argv[1]
src = Path(sys.argv[2]) if len(sys.argv) > 2 else ROOT / "gpurun_out"
prof = ROOT / "profiles"


ALIAS = {"ulk_ccl_words_kernel": "K_CclWords", "ulk_expand_kernel": "K_ExpandForeground", "ulk_front_tile_kernel": "K_FrontTile",
         "ulk_attach_vec4_kernel": "K_Attach",
         "ulk_threshold_kernel": "K_Threshold", "ulk_threshold_kernel_v4": "K_Threshold"}


def short(name):
    m = re.search(r"(K_\w+(?:<[^>]*>)?)", name)
    if m:
        return m.group(1)
    for k, v in ALIAS.items():
        if re.search(r"\b" + k + r"\b", name):
            return v
    return re.sub(r"^void |\(.*$", "", name)[:70]


for f in sorted(src.glob("launches_*.csv")):
    wl = f.stem.split("_", 1)[1]
    rows = list(csv.reader(l for l in open(f) if l.startswith('"')))
    hdr = rows[0]
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
    for r in rows[1:]:
        rec = dict(zip(hdr, r))
        v = float(rec["Metric Value"].replace(",", ""))
        u = rec["Metric Unit"]
        a = agg[short(rec["Kernel Name"])]
        if rec["Metric Name"] == "gpu__time_duration.sum":
            a[0] += 1
            a[1] += v * {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0}[u]
        else:
            a[2] += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    agg.pop("K_SynthFrame", None)      # input generation, outside the timed region
    tot = sum(a[1] for a in agg.values())
    out = prof / f"{tag}_ncu_launches_{wl}.csv"
    with open(out, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel", "launches", "total_ms", "share_of_step", "dram_bytes_per_launch", "dram_GBps"])
        for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            w.writerow([k, a[0], f"{a[1]:.4f}", f"{a[1] / tot:.4f}", int(a[2] / a[0]), f"{a[2] / 1e9 / (a[1] / 1e3):.0f}" if a[1] else ""])
    tf = prof / "ncu_traffic.json"
    traffic = json.loads(tf.read_text()) if tf.exists() else {}
    traffic[wl] = {k.split("<")[0]: int(a[2] / a[0]) for k, a in agg.items() if k.startswith("K_")}
    tf.write_text(json.dumps(traffic, indent=1, sort_keys=True))
    print("wrote", out, tf)

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]
# the report itself is too large to bring back from the GPU box: its raw page is exported there
# (full_<wl>_raw.csv); a local .ncu-rep is read through ncu when present
for f in sorted(list(src.glob("full_*.ncu-rep")) + list(src.glob("full_*_raw.csv"))):
    wl = f.stem.split("_")[1]
    if f.suffix == ".csv":
        raw = f.read_text()
    else:
        raw = subprocess.run(["ncu", "-i", str(f), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [c for c in WANT if c in idx]
    out = prof / f"{tag}_ncu_full_{wl}.csv"
    with open(out, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel"] + [f"{c} [{units[idx[c]]}]" for c in cols])
        for r in rows[2:]:
            w.writerow([short(r[idx["Kernel Name"]])] + [r[idx[c]] for c in cols])
    print("wrote", out)
