"""Deterministic random-init parameters for the update block.

The key names and shapes are those of the reference module's state_dict
(BasicSelectiveMultiUpdateBlock, core/update.py:122-139), so a
dict produced here loads into the reference module with strict=True and into
``foundationstereo_b200.update.UpdateBlockParams``.  Values are drawn per key
from a seeded CPU generator in a fixed key order, which makes them identical
in the build container and on the GPU box without shipping 65 MB of weights.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, Tuple

import torch


def update_block_param_shapes(
    corr_levels: int = 2, corr_radius: int = 4, hidden: int = 128, volume_dim: int = 28, n_gru_layers: int = 3
) -> "OrderedDict[str, Tuple[int, ...]]":
    """name -> shape; which GRUs exist and their input widths follow core/update.py:128-132."""
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()

    def conv(name, cout, cin, k):
        s[name + ".weight"] = (cout, cin, k, k)
        s[name + ".bias"] = (cout,)

    cor_planes = corr_levels * (2 * corr_radius + 1) * (volume_dim + 1)
    conv("encoder.convc1", 256, cor_planes, 1)
    conv("encoder.convc2", 256, 256, 3)
    conv("encoder.convd1", 64, 1, 7)
    conv("encoder.convd2", 64, 64, 3)
    conv("encoder.conv", hidden - 1, 64 + 256, 3)
    grus = {3: (("gru16", 2 * hidden), ("gru08", 3 * hidden), ("gru04", 3 * hidden)),
            2: (("gru08", 2 * hidden), ("gru04", 3 * hidden)),
            1: (("gru04", 2 * hidden),)}[n_gru_layers]
    for name, cin in grus:
        conv(name + ".conv0.0", cin, cin, 3)
        conv(name + ".conv1.0", cin + hidden, cin + hidden, 3)
        for gate in ("convz", "convr", "convq"):
            conv(f"{name}.small_gru.{gate}", hidden, cin + hidden, 1)
        for gate in ("convz", "convr", "convq"):
            conv(f"{name}.large_gru.{gate}", hidden, cin + hidden, 3)
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


def make_update_block_params(
    seed: int = 0,
    corr_levels: int = 2,
    corr_radius: int = 4,
    hidden: int = 128,
    volume_dim: int = 28,
    delta_scale: float = 1.0,
    n_gru_layers: int = 3,
) -> Dict[str, torch.Tensor]:
    """U(-1/sqrt(fan_in), 1/sqrt(fan_in)) weights and biases (PyTorch's default
    conv/linear range), layer-scale gammas in [0.1, 0.2] so that the EdgeNeXt
    branch of the disparity head actually contributes (the reference's init
    value 1e-6 would hide errors there).  ``delta_scale`` multiplies the last
    disparity-head conv so a free-running random-weight loop stays bounded
    (SURVEY.md H10)."""
    g = torch.Generator().manual_seed(seed)
    out: Dict[str, torch.Tensor] = {}
    shapes = update_block_param_shapes(corr_levels, corr_radius, hidden, volume_dim, n_gru_layers)
    fan_in: Dict[str, int] = {}
    for name, shp in shapes.items():
        if name.endswith(".weight"):
            fan_in[name[: -len(".weight")]] = int(torch.tensor(shp[1:]).prod())
    for name, shp in shapes.items():
        if name.endswith(".gamma"):
            out[name] = 0.1 + 0.1 * torch.rand(shp, generator=g)
            continue
        base = name.rsplit(".", 1)[0]
        bound = 1.0 / math.sqrt(fan_in[base])
        out[name] = (torch.rand(shp, generator=g) * 2 - 1) * bound
    if delta_scale != 1.0:
        out["disp_head.conv.4.weight"] *= delta_scale
        out["disp_head.conv.4.bias"] *= delta_scale
    return out
