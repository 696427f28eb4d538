"""Seeded synthetic inputs of the hot path.

Generated on the CPU from a torch.Generator so the oracle and the CUDA path
see identical bits.  TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.
"""
from __future__ import annotations

from typing import Dict

import torch


def make_inputs(
    B: int = 1,
    C: int = 128,
    H: int = 120,
    W: int = 160,
    D: int = 48,
    Cp: int = 12,
    Cg: int = 28,
    hidden: int = 128,
    seed: int = 0,
    with_volume: bool = True,
) -> Dict[str, object]:
    g = torch.Generator().manual_seed(seed)

    def randn(*s):
        return torch.randn(*s, generator=g)

    def rand(*s):
        return torch.rand(*s, generator=g)

    out: Dict[str, object] = {}
    out["fl"] = torch.relu(randn(B, C, H, W))
    out["fr"] = torch.relu(randn(B, C, H, W))
    out["pl"] = randn(B, Cp, H, W)
    out["pr"] = randn(B, Cp, H, W)
    if with_volume:
        out["volume"] = randn(B, Cg, D, H, W)
    disp = rand(B, 1, H, W) * D
    u = rand(B, 1, H, W)
    lo = -8.0 * rand(B, 1, H, W)
    hi = D + 8.0 * rand(B, 1, H, W)
    disp = torch.where(u < 0.05, lo, disp)
    disp = torch.where(u > 0.95, hi, disp)
    out["disp"] = disp
    out["coords"] = torch.arange(W, dtype=torch.float32).reshape(1, 1, W, 1).repeat(B, H, 1, 1)
    net, inp, att = [], [], []
    for i in range(3):
        h, w = _down(H, i), _down(W, i)
        net.append(torch.tanh(randn(B, hidden, h, w)))
        inp.append(torch.relu(randn(B, hidden, h, w)))
        att.append(torch.sigmoid(randn(B, 1, h, w)))
    out["net"], out["inp"], out["att"] = net, inp, att
    return out


def _down(n: int, times: int) -> int:
    """Spatial size after ``times`` stride-2 stages (conv k3 s2 p1 -> ceil(n/2)),
    which is what the reference context network produces for 1/8 and 1/16 and
    what pool2x (core/update.py:72) maps 1/4 -> 1/8 -> 1/16 to."""
    for _ in range(times):
        n = (n + 1) // 2
    return n
