#!/usr/bin/env python
"""Launches each libfsb200 kernel a few times at BASELINE configs[1] size -- the short
program `ncu --set full` is pointed at (tools/gpu_*.sh)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import foundationstereo_b200 as fsb
from foundationstereo_b200 import ops, synthetic
from foundationstereo_b200._lib import FSB_PREC_FP32

B, C, H, W, D, L, r = 1, 128, 120, 160, 104, 2, 4
B = int(sys.argv[2]) if len(sys.argv) > 2 else B
x = synthetic.hot_path_inputs(B, C, H, W, D)
d = {k: v.cuda() for k, v in x.items() if isinstance(v, torch.Tensor)}
d["volume"] = d["volume"].half()   # as bench.py hands it over: the filtered volume in the reference's autocast dtype
dx = torch.linspace(-r, r, 2 * r + 1).reshape(1, 1, -1, 1)
coords = torch.arange(W, dtype=torch.float32, device="cuda").reshape(1, 1, W, 1).repeat(B, H, 1, 1)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
Cpad = (L * (2 * r + 1) * 29 + 63) // 64 * 64
out16 = torch.zeros(B, H, W, Cpad, dtype=torch.float16, device="cuda")
cls_logits = torch.randn(B, D, H, W, device="cuda")
spx_logits = torch.randn(B, 9, 4 * H, 4 * W, device="cuda")
for _ in range(reps):
    ops.disparity_regression(cls_logits, True)
    ops.context_upsample(d["disp"], spx_logits, True, 4.0)
    ops.combined_volume(d["fl"], d["fr"], d["pl"], d["pr"], D, 8)
    geo = fsb.Combined_Geo_Encoding_Volume(d["fl"], d["fr"], d["volume"], num_levels=L, dx=dx)
    out = geo(d["disp"], coords)
    geo.lookup_nhwc16(d["disp"], coords, out16)
    band = fsb.Combined_Geo_Encoding_Volume(d["fl"], d["fr"], d["volume"], num_levels=L, dx=dx, corr_band=D)
    band.lookup_nhwc16(d["disp"], coords, out16)
torch.cuda.synchronize()
print("ok", out.shape, fsb.launch_count())
