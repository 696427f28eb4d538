"""Seeded synthetic inputs and random-init update-block weights of the named shapes
(SURVEY.md 8(d)): used by bench.py and __graft_entry__.smoke(), where no dataset or
checkpoint can be downloaded.  Generated on the CPU so every rank and every run sees the
same bits."""
from __future__ import annotations

import math
from typing import Dict

import torch


def down2(n: int, times: int) -> int:
    """size after `times` stride-2 stages (k3 s2 p1 -> ceil(n/2)): 1/4 -> 1/8 -> 1/16."""
    for _ in range(times):
        n = (n + 1) // 2
    return n


def hot_path_inputs(B: int, C: int, H: int, W: int, D: int, Cp: int = 12, Cg: int = 28, hidden: int = 128,
                    seed: int = 0) -> Dict[str, object]:
    g = torch.Generator().manual_seed(seed)
    rn = lambda *s: torch.randn(*s, generator=g)
    ru = lambda *s: torch.rand(*s, generator=g)
    x: Dict[str, object] = {
        "fl": torch.relu(rn(B, C, H, W)), "fr": torch.relu(rn(B, C, H, W)),
        "pl": rn(B, Cp, H, W), "pr": rn(B, Cp, H, W),
        "volume": rn(B, Cg, D, H, W),
    }
    disp = ru(B, 1, H, W) * D
    u = ru(B, 1, H, W)
    disp = torch.where(u < 0.05, -8.0 * ru(B, 1, H, W), disp)
    disp = torch.where(u > 0.95, D + 8.0 * ru(B, 1, H, W), disp)
    x["disp"] = disp
    x["net"] = [torch.tanh(rn(B, hidden, down2(H, i), down2(W, i))) for i in range(3)]
    x["inp"] = [torch.relu(rn(B, hidden, down2(H, i), down2(W, i))) for i in range(3)]
    x["att"] = [torch.sigmoid(rn(B, 1, down2(H, i), down2(W, i))) for i in range(3)]
    return x


def update_block_shapes(corr_levels: int = 2, corr_radius: int = 4, hidden: int = 128, volume_dim: int = 28):
    """state_dict key -> shape of BasicSelectiveMultiUpdateBlock with n_gru_layers == 3
    (core/update.py:122-139)."""
    s = {}

    def conv(name, cout, cin, k):
        s[name + ".weight"] = (cout, cin, k, k)
        s[name + ".bias"] = (cout,)

    planes = corr_levels * (2 * corr_radius + 1) * (volume_dim + 1)
    conv("encoder.convc1", 256, planes, 1)
    conv("encoder.convc2", 256, 256, 3)
    conv("encoder.convd1", 64, 1, 7)
    conv("encoder.convd2", 64, 64, 3)
    conv("encoder.conv", hidden - 1, 64 + 256, 3)
    for name, cin in (("gru16", 2 * hidden), ("gru08", 3 * hidden), ("gru04", 3 * hidden)):
        conv(name + ".conv0.0", cin, cin, 3)
        conv(name + ".conv1.0", cin + hidden, cin + hidden, 3)
        for gru, k in (("small_gru", 1), ("large_gru", 3)):
            for gate in ("convz", "convr", "convq"):
                conv(f"{name}.{gru}.{gate}", hidden, cin + hidden, k)
    conv("disp_head.conv.0", hidden, hidden, 3)
    for i in (2, 3):
        pre = f"disp_head.conv.{i}"
        s[pre + ".gamma"] = (hidden,)
        s[pre + ".dwconv.weight"] = (hidden, 1, 7, 7)
        s[pre + ".dwconv.bias"] = (hidden,)
        s[pre + ".pwconv1.weight"] = (4 * hidden, hidden)
        s[pre + ".pwconv1.bias"] = (4 * hidden,)
        s[pre + ".pwconv2.weight"] = (hidden, 4 * hidden)
        s[pre + ".pwconv2.bias"] = (hidden,)
    conv("disp_head.conv.4", 1, hidden, 3)
    conv("mask.0", 64, hidden, 3)
    conv("mask.2", 32, 64, 3)
    return s


def random_update_block_state(seed: int = 0, corr_levels: int = 2, corr_radius: int = 4, hidden: int = 128,
                              volume_dim: int = 28, delta_scale: float = 1.0) -> Dict[str, torch.Tensor]:
    """PyTorch's default conv/linear init range (U(-1/sqrt(fan_in), 1/sqrt(fan_in))), untouched by default:
    32 free-running iterations stay bounded and within 0.005 px of the fp32 oracle at full size
    (profiles/r02a_fullsize_parity.json).  `delta_scale` multiplies the last disparity-head conv for tests."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for name, shp in update_block_shapes(corr_levels, corr_radius, hidden, volume_dim).items():
        if name.endswith(".gamma"):
            out[name] = 0.1 + 0.1 * torch.rand(shp, generator=g)
            continue
        wshape = shp if name.endswith(".weight") else update_block_shapes(corr_levels, corr_radius, hidden, volume_dim)[
            name[: -len(".bias")] + ".weight"]
        bound = 1.0 / math.sqrt(math.prod(wshape[1:]))
        out[name] = (torch.rand(shp, generator=g) * 2 - 1) * bound
    out["disp_head.conv.4.weight"] *= delta_scale
    out["disp_head.conv.4.bias"] *= delta_scale
    return out


def update_block_flops(state: Dict[str, torch.Tensor], B: int, H: int, W: int) -> float:
    """Algorithmic FLOPs of one BasicSelectiveMultiUpdateBlock.forward (core/update.py:141-159) on the
    GEMM-shaped layers (every Conv2d / Linear except the 1-channel and depthwise 7x7 convolutions):
    2 * pixels * Cout * Cin * k * k, pixels = those of the pyramid level the layer runs on."""
    px = [B * down2(H, i) * down2(W, i) for i in range(3)]
    level = {"gru04": 0, "gru08": 1, "gru16": 2}
    total = 0.0
    for name, w in state.items():
        if not name.endswith(".weight") or name.endswith("dwconv.weight") or name == "encoder.convd1.weight":
            continue
        n = px[level.get(name.split(".")[0], 0)]
        total += 2.0 * n * w.numel()
    return total
