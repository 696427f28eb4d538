#!/usr/bin/env python
"""Times fsb_corr_pyramid alone (8 pairs at BASELINE configs[1] size by default), L2 flushed before each launch."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foundationstereo_b200 import ops
from foundationstereo_b200._lib import FSB_PREC_TF32X3, FSB_PREC_TF32

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
C, H, W, L = 128, 120, 160, 2
g = torch.Generator().manual_seed(0)
f1 = torch.relu(torch.randn(B, C, H, W, generator=g)).cuda()
f2 = torch.relu(torch.randn(B, C, H, W, generator=g)).cuda()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
alg = 8 * B * H * W * C + 4 * B * H * W * sum(W >> l for l in range(L))
for prec, name in ((FSB_PREC_TF32X3, "tf32x3"), (FSB_PREC_TF32, "tf32")):
    levels = [torch.empty((B * H * W, W >> l), device="cuda") for l in range(L)]
    for _ in range(3):
        ops.corr_pyramid_into(f1, f2, levels, prec)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for a, b in ev:
        flush.zero_()
        a.record()
        ops.corr_pyramid_into(f1, f2, levels, prec)
        b.record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
    print(f"{name} dbg={os.environ.get('FSB_CORR_DBG', '0')}: {ms * 1e3:.1f} us  {alg / ms / 1e6:.0f} GB/s")
