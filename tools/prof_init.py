#!/usr/bin/env python
"""Device times of the once-per-pair kernels (1, 2, 2b) and the lookup (3) at BASELINE configs[1] size,
with and without an L2 flush before each launch.   python tools/prof_init.py [pairs]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import foundationstereo_b200 as fsb
from foundationstereo_b200 import ops, synthetic
from foundationstereo_b200._lib import FSB_PREC_FP32, FSB_PREC_TF32, FSB_PREC_TF32X3

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
C, H, W, D, L, r = 128, 120, 160, 104, 2, 4
x = synthetic.hot_path_inputs(B, C, H, W, D)
d = {k: v.cuda() for k, v in x.items() if isinstance(v, torch.Tensor)}
dx = torch.linspace(-r, r, 2 * r + 1).reshape(1, 1, -1, 1)
coords = torch.arange(W, dtype=torch.float32, device="cuda").reshape(1, 1, W, 1).repeat(B, H, 1, 1)
geo = fsb.Combined_Geo_Encoding_Volume(d["fl"], d["fr"], d["volume"], num_levels=L, dx=dx)
buf = torch.empty(B, H, W, 576, dtype=torch.float16, device="cuda")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
fns = {
    "volume": lambda: ops.combined_volume(d["fl"], d["fr"], d["pl"], d["pr"], D, 8),
    "corr fp32": lambda: ops.corr_pyramid(d["fl"], d["fr"], L, FSB_PREC_FP32),
    "corr tf32x3": lambda: ops.corr_pyramid(d["fl"], d["fr"], L, FSB_PREC_TF32X3),
    "corr tf32": lambda: ops.corr_pyramid(d["fl"], d["fr"], L, FSB_PREC_TF32),
    "geo_pyramid": lambda: ops.geo_pyramid(d["volume"], L),
    "lookup nhwc16": lambda: ops.geo_lookup_nhwc16(geo.geo_levels, geo.corr_levels, d["disp"], coords, r, buf),
    "lookup fp32": lambda: geo(d["disp"], coords),
}
for name, fn in fns.items():
    for _ in range(3):
        fn()
    res = []
    for do_flush in (False, True):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
        for a, b in ev:
            if do_flush:
                flush.zero_()
            a.record()
            fn()
            b.record()
        torch.cuda.synchronize()
        res.append(1e3 * sum(a.elapsed_time(b) for a, b in ev) / len(ev))
    print(f"{name:16s} pairs {B}: warm-L2 {res[0]:8.1f} us   after 512 MiB flush {res[1]:8.1f} us")
