#!/usr/bin/env python
"""ncu source page of one kernel -> executed warp instructions by opcode + hottest stall lines.
    python tools/ncu_src.py <rep> <kernel-regex> [warps]"""
import collections
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx],
                     capture_output=True, text=True).stdout
allrows = list(csv.reader(raw.splitlines()))
starts = [i for i, r in enumerate(allrows) if r and r[0] == "Kernel Name"]
want = int(sys.argv[3]) if len(sys.argv) > 3 else 0  # n-th block of the page (each launch appears as SASS block(s))
sub = [b for b in starts if rx.replace("\\", "") in allrows[b][1] or True]
b0 = sub[want]
b1 = starts[starts.index(b0) + 1] if starts.index(b0) + 1 < len(starts) else len(allrows)
rows = allrows[b0:b1]
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[2:] if len(r) == len(hdr) and r[0].startswith("0x")]
E = ix["Instructions Executed"]
S = ix["Warp Stall Sampling (All Samples)"]
tot = sum(int(r[E]) for r in data)
print(rows[0][1][:100])
print("executed warp instructions", tot, "static", len(data))
ops = collections.Counter()
for r in data:
    t = r[ix["Source"]].strip().split()
    op = t[1] if t[0].startswith("@") else t[0]
    ops[op.split(".")[0]] += int(r[E])
print(" ".join(f"{k}:{100 * v / tot:.1f}%" for k, v in ops.most_common(18)))
stot = sum(int(r[S]) for r in data)
for r in sorted(data, key=lambda r: -int(r[S]))[:14]:
    print(f"{100 * int(r[S]) / stot:5.1f}%  exec {int(r[E]):9d}  {r[ix['Source']].strip()[:80]}")
