#!/usr/bin/env python
"""Times the post kernels alone (8 pairs at BASELINE configs[1] size), L2 flushed before each launch."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foundationstereo_b200 import ops
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
D, H, W = 104, 120, 160
logits = torch.randn(B, D, H, W, device="cuda"); spx = torch.randn(B, 9, 4 * H, 4 * W, device="cuda")
disp = torch.rand(B, 1, H, W, device="cuda") * D
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
def t(fn, nbytes, name):
    for _ in range(3): fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for a, b in ev:
        flush.zero_(); a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
    print(f"{name}: {ms * 1e3:.1f} us  {nbytes / ms / 1e6:.0f} GB/s")
px = B * H * W
t(lambda: ops.disparity_regression(logits, True), px * (D + 1) * 4, "softmax_disparity_regression")
t(lambda: ops.disparity_regression(torch.softmax(logits, 1), False) if False else ops.disparity_regression(logits, False), px * (D + 1) * 4, "disparity_regression")
t(lambda: ops.context_upsample(disp, spx, True, 4.0), px * (4 + 160 * 4), "softmax_context_upsample")
t(lambda: ops.context_upsample(disp, spx, False, 1.0), px * (4 + 160 * 4), "context_upsample")
