#!/usr/bin/env python
"""ncu source page (sass,cuda view; needs -lineinfo + --import-source on): stall samples per SOURCE LINE of one kernel,
with the split into stall reasons.
    python tools/ncu_lines.py <rep> <kernel-regex> [top]"""
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda", "--kernel-name",
                      "regex:" + rx], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
agg, reasons = {}, {}
fname = hdr = first = cur = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        first = first or r[1]
        cur = r[1]
        continue
    if r[0] == "Line No":
        hdr = {}
        for i, h in enumerate(r):
            hdr.setdefault(h, i)
        continue
    if hdr is None or cur != first or not r[0].isdigit():   # only the per-line summary rows
        continue
    try:
        s = int(r[hdr["Warp Stall Sampling (All Samples)"]])
        e = int(r[hdr["Instructions Executed"]])
    except (ValueError, IndexError):
        continue
    key = (fname, int(r[0]), r[1].strip()[:100])
    a = agg.setdefault(key, [0, 0, {}])
    a[0] += s
    a[1] += e
    for h, i in hdr.items():
        if h.startswith("stall_") and "Not Issued" not in h and i < len(r) and r[i].isdigit() and int(r[i]):
            a[2][h[6:]] = a[2].get(h[6:], 0) + int(r[i])
            reasons[h[6:]] = reasons.get(h[6:], 0) + int(r[i])
tot = sum(v[0] for v in agg.values()) or 1
print(first[:100], "samples", tot)
print("reasons:", " ".join(f"{k}:{100 * v / tot:.0f}%" for k, v in sorted(reasons.items(), key=lambda kv: -kv[1])[:8]))
for (f, ln, src), (s, e, rs) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    why = ",".join(f"{k}{100 * v // max(s, 1)}" for k, v in sorted(rs.items(), key=lambda kv: -kv[1])[:2])
    print(f"{100 * s / tot:5.1f}%  exec {e:8d}  {f}:{ln}  [{why}]  {src}")
