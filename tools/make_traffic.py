#!/usr/bin/env python
"""profiles/traffic.json from two `ncu --set full` captures (8 pairs per launch, as bench.py's default):
    python tools/make_traffic.py <update-step capture .ncu-rep> <kernels 1-3 capture .ncu-rep> [<lookup_convc1 capture>]
DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) per launch; for the update block the sum over the
tensor-core launches of ONE step (what bench.py's roofline.achieved is defined over)."""
import csv
import json
import os
import subprocess
import sys


def rows(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(raw.splitlines()))
    hdr, units = r[0], r[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def to_bytes(v, u):
        v = float(v.replace(",", ""))
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    out = []
    for x in r[2:]:
        b = sum(to_bytes(x[ix[m]], units[ix[m]]) for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        out.append((x[ix["Kernel Name"]], b, float(x[ix["gpu__time_duration.sum"]].replace(",", ""))))
    return out


step, k123 = rows(sys.argv[1]), rows(sys.argv[2])
t = {}
tensor = [(n, b) for n, b, _ in step if "conv_umma" in n or "conv_halo" in n or "edgenext_mlp" in n]
t["conv_umma_kernel"] = sum(b for _, b in tensor)
t["conv_umma_kernel_launches"] = len(tensor)
names = {"volume_kernel": "volume_gwc_concat", "corr_umma": "corr_pyramid", "corr_tma": "corr_pyramid",
         "geo_pyramid": "geo_pyramid", "lookup_nhwc": "geo_lookup", "lookup_nchw": "geo_lookup_fp32_nchw",
         "disp_regression": "softmax_disparity_regression", "context_upsample": "softmax_context_upsample"}
for n, b, _ in k123:
    for k, v in names.items():
        if k in n:
            if v == "corr_pyramid" and v in t:
                t["corr_pyramid_band"] = b  # tools/prof_kernels.py launches the full form first, then the band form
            elif v == "geo_lookup" and v in t:
                t["geo_lookup_band"] = b
            elif v not in t:
                t[v] = b  # first launch of each kernel
if len(sys.argv) > 3:   # optional third capture: the lookup fused into encoder.convc1 (tools/lookconv_exp.py)
    for n, b, _ in rows(sys.argv[3]):
        if "lookup_convc1" in n:
            t.setdefault("lookup_convc1", b)
t["note"] = "bytes per launch at 8 pairs (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum); conv_umma_kernel = " \
            "sum over the tensor-core launches of one update step"
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
with open(os.path.join(root, "profiles", "traffic.json"), "w") as fh:
    json.dump(t, fh, indent=1)
print(json.dumps(t, indent=1))
