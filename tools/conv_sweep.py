#!/usr/bin/env python
"""Timing sweep of fsb_conv2d_nhwc over synthetic layers (ReLU epilogue unless noted):
    python tools/conv_sweep.py [B H W]
Each line: name, layout, block_n, us, TFLOP/s.  Used to separate main-loop cost from epilogue cost."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foundationstereo_b200 import _lib
from foundationstereo_b200.conv import ConvLaunch, pack_rows, pack_segment

B, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (4, 120, 160)
dev = "cuda"
dt = torch.float16


def run(name, cins, cout, k, bn, layout, reps=20, stages=0):
    srcs = [torch.randn(B, H, W, c, device=dev, dtype=dt) for c in cins]
    w = torch.randn(cout, sum(cins), k, k, device=dev) * 0.02
    mat = pack_segment(w, [(c, c) for c in cins], layout)
    rows = (cout + bn - 1) // bn * bn
    wp = pack_rows([mat], rows).to(dt).contiguous()
    bias = torch.zeros(1, rows, device=dev)
    out = torch.empty(B, H, W, rows, device=dev, dtype=dt)
    L = ConvLaunch(srcs=srcs, segs=[(k, list(range(len(cins))))], weight=wp, bias=bias, cout=cout, block_n=bn,
                   epilogue=_lib.FSB_EPI_RELU, out=out, layout=layout, stages=stages, name=name)
    for _ in range(3):
        L()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        L()
    e1.record()
    torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / reps
    print(f"{name:28s} lay={layout} bn={bn:3d} st={stages} {us:8.1f} us {L.flops / us * 1e-6:8.1f} TFLOP/s", flush=True)


TAP, HALO = _lib.FSB_CONV_LAYOUT_TAP, _lib.FSB_CONV_LAYOUT_HALO
only = os.environ.get("SWEEP_ONLY")  # substring filter on "name bn"
CFG = [
    ("3x3 512->128", [128, 384], 128, 3, 64, TAP), ("3x3 512->128", [128, 384], 128, 3, 128, TAP),
    ("3x3 512->128", [128, 384], 128, 3, 128, HALO),
    ("3x3 512->512", [384, 128], 512, 3, 128, TAP), ("3x3 512->512", [384, 128], 512, 3, 256, TAP),
    ("3x3 512->512", [384, 128], 512, 3, 256, HALO),
    ("3x3 384->384", [128, 128, 128], 384, 3, 128, TAP), ("3x3 384->384", [128, 128, 128], 384, 3, 192, TAP),
    ("3x3 384->384", [128, 128, 128], 384, 3, 192, HALO),
    ("3x3 512->256", [512], 256, 3, 128, TAP), ("3x3 512->256", [512], 256, 3, 256, TAP),
    ("3x3 256->256", [256], 256, 3, 128, TAP), ("3x3 256->256", [256], 256, 3, 256, TAP),
    ("3x3 320->128", [256, 64], 128, 3, 64, TAP), ("3x3 320->128", [256, 64], 128, 3, 128, TAP),
    ("3x3 128->128", [128], 128, 3, 64, TAP), ("3x3 128->128", [128], 128, 3, 128, TAP),
    ("3x3 128->64", [128], 64, 3, 64, TAP), ("3x3 64->64", [64], 64, 3, 64, TAP),
    ("1x1 576->256", [576], 256, 1, 128, TAP), ("1x1 576->256", [576], 256, 1, 256, TAP),
    ("1x1 128->512", [128], 512, 1, 128, TAP), ("1x1 128->512", [128], 512, 1, 256, TAP),
    ("1x1 512->128", [512], 128, 1, 64, TAP), ("1x1 512->128", [512], 128, 1, 128, TAP),
]
for name, cins, cout, k, bn, lay in CFG:
    if only and only not in f"{name} bn{bn} lay{lay}":
        continue
    run(name, cins, cout, k, bn, lay, reps=int(os.environ.get("SWEEP_REPS", "20")))
