#!/usr/bin/env python
"""Per-kernel counts of the SASS mnemonics that prove a Blackwell-native kernel (B200_PROFILING.md): tcgen05.mma ->
UTC*MMA, tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG/UBLKCP, packed fp32 -> FFMA2; HMMA would be the legacy
mma.sync path.  Reads the in-tree libfsb200.so with cuobjdump (no GPU needed).

    python tools/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "foundationstereo_b200", "libfsb200.so")
COLS = ["UTCHMMA", "UTCQMMA", "UTCMMA_OTHER", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "FFMA2", "FFMA", "HMMA",
        "MUFU", "LDGSTS"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kern = None
    counts = collections.OrderedDict()
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            kern = kern.replace("(anonymous namespace)::", "")
            kern = re.sub(r"\(.*", "", kern).replace("void ", "").replace("fsb::", "")
            counts[kern] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if not m or kern is None:
            continue
        op = m.group(1)
        c = counts[kern]
        if op.startswith("UTC") and "MMA" in op:
            c["UTCHMMA" if op.startswith("UTCHMMA") else "UTCQMMA" if op.startswith("UTCQMMA") else "UTCMMA_OTHER"] += 1
        elif op in ("LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "FFMA2", "FFMA", "MUFU", "LDGSTS"):
            c[op] += 1
        elif op.startswith("HMMA"):
            c["HMMA"] += 1
        c["total"] += 1
    arch = subprocess.run(["cuobjdump", "--list-elf", LIB], capture_output=True, text=True).stdout.strip().splitlines()
    print("# libfsb200.so: " + "; ".join(a.split()[-1] for a in arch))
    print("# static SASS instruction counts per kernel (cuobjdump -sass); UTC*MMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st,")
    print("# UTMALDG/UTMASTG = TMA tensor load/store, SYNCS = mbarrier ops, FFMA2 = packed fp32 FMA; HMMA (mma.sync) must be 0")
    print(f"{'kernel':66s} " + " ".join(f"{c:>8s}" for c in COLS) + f" {'total':>8s}")
    for k, c in counts.items():
        print(f"{k[:66]:66s} " + " ".join(f"{c.get(col, 0):8d}" for col in COLS) + f" {c['total']:8d}")
    assert all(c.get("HMMA", 0) == 0 for c in counts.values()), "legacy mma.sync in the library"


if __name__ == "__main__":
    main()
