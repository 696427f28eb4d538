"""Update block of the refinement loop: BasicSelectiveMultiUpdateBlock.forward
(core/update.py:141-159) as a functional engine over the reference's state_dict.

INTERIM (round 1, first GPU pass): the convolutions below are still issued through
torch (cuDNN, fp16 operands / fp32 accumulate, the same arithmetic the reference runs
under its autocast region, core/foundation_stereo.py:239).  They are being replaced by
the tcgen05 implicit-GEMM kernel of libfsb200.so (fsb_conv2d_nhwc); this file keeps the
module contract (parameter names, in-place ``net`` mutation, return order) so the switch
does not move the boundary.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


class SelectiveUpdateBlock:
    """Holds fp16 copies of the update-block weights (keys = reference state_dict keys,
    SURVEY.md 8(b)) and runs one refinement step on CUDA tensors."""

    def __init__(self, state_dict: Dict[str, Tensor], device: torch.device | str = "cuda",
                 dtype: torch.dtype = torch.float16):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("SelectiveUpdateBlock runs on CUDA only (no CPU path)")
        self.dtype = dtype
        self.p: Dict[str, Tensor] = {}
        for k, v in state_dict.items():
            t = v.detach().to(self.device, dtype)
            if t.dim() == 4:
                t = t.contiguous(memory_format=torch.channels_last)
            self.p[k] = t
        self.n_gru_layers = 3
        assert "gru16.conv0.0.weight" in self.p and "gru08.conv0.0.weight" in self.p, \
            "only the n_gru_layers == 3 configuration of the reference is supported"

    # ---- building blocks -----------------------------------------------------------
    def _conv(self, name: str, x: Tensor, pad: int, groups: int = 1) -> Tensor:
        return F.conv2d(x, self.p[name + ".weight"], self.p[name + ".bias"], padding=pad, groups=groups)

    def _raft_gru(self, pre: str, h: Tensor, x: Tensor, hx: Tensor) -> Tensor:
        pad = self.p[pre + ".convz.weight"].shape[-1] // 2
        z = torch.sigmoid(self._conv(pre + ".convz", hx, pad))
        r = torch.sigmoid(self._conv(pre + ".convr", hx, pad))
        q = torch.tanh(self._conv(pre + ".convq", torch.cat([r * h, x], 1), pad))
        return (1 - z) * h + z * q

    def _selective_gru(self, pre: str, att: Tensor, h: Tensor, *xs: Tensor) -> Tensor:
        x = F.relu(self._conv(pre + ".conv0.0", torch.cat(xs, 1), 1))
        hx = F.relu(self._conv(pre + ".conv1.0", torch.cat([x, h], 1), 1))
        return self._raft_gru(pre + ".small_gru", h, x, hx) * att + self._raft_gru(pre + ".large_gru", h, x, hx) * (1 - att)

    def _encoder(self, disp: Tensor, corr: Tensor) -> Tensor:
        c = F.relu(self._conv("encoder.convc1", corr, 0))
        c = F.relu(self._conv("encoder.convc2", c, 1))
        d = F.relu(self._conv("encoder.convd1", disp, 3))
        d = F.relu(self._conv("encoder.convd2", d, 1))
        o = F.relu(self._conv("encoder.conv", torch.cat([c, d], 1), 1))
        return torch.cat([o, disp], 1)

    def _edgenext(self, pre: str, x: Tensor) -> Tensor:
        y = self._conv(pre + ".dwconv", x, 3, groups=x.shape[1]).permute(0, 2, 3, 1)
        y = F.gelu(F.linear(y, self.p[pre + ".pwconv1.weight"], self.p[pre + ".pwconv1.bias"]))
        y = F.linear(y, self.p[pre + ".pwconv2.weight"], self.p[pre + ".pwconv2.bias"])
        return x + (self.p[pre + ".gamma"] * y).permute(0, 3, 1, 2)

    # ---- one step -------------------------------------------------------------------
    @torch.no_grad()
    def forward(self, net: List[Tensor], inp: Sequence[Tensor], corr: Tensor, disp: Tensor,
                att: Sequence[Tensor]) -> Tuple[List[Tensor], Tensor, Tensor]:
        """Same contract as core/update.py:141-159: mutates and returns ``net``; returns
        (net, mask, delta_disp) with fp32 outputs."""
        dt = self.dtype
        n = [t.to(dt) for t in net]
        i = [t.to(dt) for t in inp]
        a = [t.to(dt) for t in att]
        n[2] = self._selective_gru("gru16", a[2], n[2], i[2], F.avg_pool2d(n[1], 3, stride=2, padding=1))
        up = F.interpolate(n[2], n[1].shape[2:], mode="bilinear", align_corners=True)
        n[1] = self._selective_gru("gru08", a[1], n[1], i[1], F.avg_pool2d(n[0], 3, stride=2, padding=1), up)
        motion = torch.cat([i[0], self._encoder(disp.to(dt), corr.to(dt))], 1)
        up = F.interpolate(n[1], n[0].shape[2:], mode="bilinear", align_corners=True)
        n[0] = self._selective_gru("gru04", a[0], n[0], motion, up)
        y = F.relu(self._conv("disp_head.conv.0", n[0], 1))
        y = self._edgenext("disp_head.conv.2", y)
        y = self._edgenext("disp_head.conv.3", y)
        delta = self._conv("disp_head.conv.4", y, 1)
        m = F.relu(self._conv("mask.0", n[0], 1))
        m = F.relu(self._conv("mask.2", m, 1))
        for k in range(3):
            net[k] = n[k].float()
        return net, 0.25 * m.float(), delta.float()

    __call__ = forward
