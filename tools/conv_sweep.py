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


def run_gate(kind, bn, reps=20):
    """GRU gate launches of gru04 (hidden 128, x 384 channels) with their real epilogues."""
    from foundationstereo_b200.conv import interleave_zr
    hd, cin = 128, 384
    new = lambda c, d=dt: torch.randn(B, H, W, c, device=dev).to(d)
    h32, att = new(hd, torch.float32), torch.rand(B, H, W, device=dev)
    z_l, z_s, rh_l, rh_s, h16, x16, hx16 = new(hd), new(hd), new(hd), new(hd), new(hd), new(cin), new(cin + hd)
    if kind == "zr":
        wl = interleave_zr(torch.randn(hd, cin + hd, 3, 3, device=dev) * 0.02, torch.randn(hd, cin + hd, 3, 3, device=dev) * 0.02, bn)
        ws = interleave_zr(torch.randn(hd, cin + hd, 1, 1, device=dev) * 0.02, torch.randn(hd, cin + hd, 1, 1, device=dev) * 0.02, bn)
        mats = [pack_segment(wl, [(cin + hd, cin + hd)], TAP), pack_segment(ws, [(cin + hd, cin + hd)], TAP)]
        wp = pack_rows(mats, 2 * hd).to(dt).contiguous()
        L = ConvLaunch(srcs=[hx16], segs=[(3, [0]), (1, [0])], weight=wp, bias=torch.zeros(2, 2 * hd, device=dev), cout=2 * hd,
                       block_n=bn, epilogue=_lib.FSB_EPI_GATE_ZR, h32=h32, z_l=z_l, z_s=z_s, rh_l=rh_l, rh_s=rh_s, hidden=hd)
    else:
        rhx = [(hd, hd), (cin, cin)]
        mats = [pack_segment(torch.randn(hd, cin + hd, 3, 3, device=dev) * 0.02, rhx, TAP),
                pack_segment(torch.randn(hd, cin + hd, 1, 1, device=dev) * 0.02, rhx, TAP)]
        wp = pack_rows(mats, hd).to(dt).contiguous()
        L = ConvLaunch(srcs=[rh_l, rh_s, x16], segs=[(3, [0, 2]), (1, [1, 2])], weight=wp, bias=torch.zeros(2, hd, device=dev),
                       cout=hd, block_n=bn, epilogue=_lib.FSB_EPI_GATE_Q, h32=h32, att=att, z_l=z_l, z_s=z_s,
                       h32_out=h32, h16_out=h16, hidden=hd)
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
    print(f"{'gate ' + kind:28s} lay=0 bn={bn:3d}      {us:8.1f} us {L.flops / us * 1e-6:8.1f} TFLOP/s", flush=True)


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
if not only or "gate" in only:
    for kind, bn in (("zr", 64), ("zr", 128), ("zr", 256), ("q", 32), ("q", 64), ("q", 128)):
        run_gate(kind, bn)
for name, cins, cout, k, bn, lay in CFG:
    if only and only not in f"{name} bn{bn} lay{lay}":
        continue
    run(name, cins, cout, k, bn, lay, reps=int(os.environ.get("SWEEP_REPS", "20")))
