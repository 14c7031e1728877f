#!/usr/bin/env python
"""Summary of an `ncu --page source --csv` export (SASS view): instructions per warp, stall mix,
the most sampled instructions.  usage: ncu_src_summary.py file.csv [instance] [top]"""
import collections
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 14
    hdr, inst, cur, names = None, [], None, []
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = []; inst.append(cur); names.append(r[1]); continue
        if r and r[0] == "Address":
            hdr = r; continue
        if cur is not None and len(r) > 6:
            cur.append(r)
    sizes = [int(c[0][5]) for c in inst]
    if which < 0:
        which = max(range(len(inst)), key=lambda i: sizes[i])
    c = inst[which]
    warps = int(c[0][5])
    print(names[which][:100], "| instances", len(inst), "| warps", warps, "| SASS", len(c),
          "| instr/warp %.1f" % (sum(int(r[5]) for r in c) / warps))
    stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[4]) for r in c)
    agg = collections.Counter()
    for r in c:
        for i in stall:
            if r[i]:
                agg[hdr[i][6:]] += int(r[i])
    print({k: round(100 * v / tot, 1) for k, v in agg.most_common(8)})
    idx = sorted(range(len(c)), key=lambda i: -int(c[i][4]))[:top]
    for i in sorted(idx):
        r = c[i]
        t2 = sorted(((int(r[j]) if r[j] else 0, hdr[j][6:]) for j in stall), reverse=True)[:1]
        print(f"{i:4d} {100 * int(r[4]) / tot:5.1f}%  x{int(r[5]) / warps:5.2f}  {r[1].strip()[:64]:64s} {t2[0][1]}")


if __name__ == "__main__":
    main()
