"""Drop-in replacements for the volume builders of the reference's core/submodule.py.

Same names, argument meaning and result (shape, dtype, contiguity) as
core/submodule.py:399-412 (build_gwc_volume, with groupwise_correlation :388-397)
and core/submodule.py:416-427 (build_concat_volume); the work is done by kernel 1
of libfsb200.so (csrc/volume.cu).
"""
from __future__ import annotations

import torch

from . import ops


def build_gwc_volume(refimg_fea: torch.Tensor, targetimg_fea: torch.Tensor, maxdisp: int, num_groups: int,
                     stride: int = 1) -> torch.Tensor:
    """(B,C,H,W) x2 -> (B,num_groups,maxdisp,H,W), dtype of ``refimg_fea``; fp32 math.
    ``stride`` is accepted and ignored exactly like the reference (core/submodule.py:399)."""
    assert refimg_fea.shape[1] % num_groups == 0, f"C={refimg_fea.shape[1]} not divisible by num_groups={num_groups}"
    if targetimg_fea.dtype != refimg_fea.dtype:
        targetimg_fea = targetimg_fea.to(refimg_fea.dtype)
    return ops.gwc_volume(refimg_fea, targetimg_fea, int(maxdisp), int(num_groups))


def build_concat_volume(refimg_fea: torch.Tensor, targetimg_fea: torch.Tensor, maxdisp: int) -> torch.Tensor:
    """(B,C,H,W) x2 -> (B,2C,maxdisp,H,W), dtype of ``refimg_fea`` (core/submodule.py:416)."""
    if targetimg_fea.dtype != refimg_fea.dtype:
        targetimg_fea = targetimg_fea.to(refimg_fea.dtype)
    return ops.concat_volume(refimg_fea, targetimg_fea, int(maxdisp))


def build_combined_volume(fl: torch.Tensor, fr: torch.Tensor, pl: torch.Tensor, pr: torch.Tensor, maxdisp: int,
                          num_groups: int) -> torch.Tensor:
    """torch.cat([build_gwc_volume(fl,fr), build_concat_volume(pl,pr)], 1) in one launch and
    without the extra copy (core/foundation_stereo.py:207-212).  Opt-in: the
    unchanged reference caller uses the two functions above."""
    return ops.combined_volume(fl, fr, pl, pr, int(maxdisp), int(num_groups))
