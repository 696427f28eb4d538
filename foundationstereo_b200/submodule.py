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


# ---- the steps either side of the refinement loop -------------------------
def disparity_regression(x: torch.Tensor, maxdisp: int) -> torch.Tensor:
    """(B,maxdisp,H,W) probabilities -> (B,1,H,W) expected disparity, dtype of ``x``
    (core/submodule.py:431-435, called at core/foundation_stereo.py:220)."""
    assert len(x.shape) == 4
    assert x.shape[1] == maxdisp, f"x has {x.shape[1]} disparities, maxdisp={maxdisp}"
    return ops.disparity_regression(x, False)


def softmax_disparity_regression(logits: torch.Tensor) -> torch.Tensor:
    """disparity_regression(F.softmax(logits, 1), D) in one pass over the logits
    (core/foundation_stereo.py:218-220): the probability tensor is never written.  fp32 result.
    Opt-in: needs the caller to hand over the classifier output instead of the softmax."""
    return ops.disparity_regression(logits, True)


def context_upsample(disp_low: torch.Tensor, up_weights: torch.Tensor) -> torch.Tensor:
    """Convex 4x upsampling (core/submodule.py:456-468): (b,1,h,w) x (b,9,4h,4w) -> (b,4h,4w)."""
    out = ops.context_upsample(disp_low, up_weights, False, 1.0)
    res = torch.result_type(disp_low, up_weights)
    return out if res == torch.float32 else out.to(res)


def softmax_context_upsample(disp: torch.Tensor, spx_logits: torch.Tensor, scale: float = 4.0) -> torch.Tensor:
    """context_upsample(disp*scale, F.softmax(spx_logits, 1)).unsqueeze(1) in one launch
    (core/foundation_stereo.py:188-189): (b,1,h,w), (b,9,4h,4w) -> (b,1,4h,4w) fp32.  Opt-in."""
    return ops.context_upsample(disp, spx_logits, True, float(scale)).unsqueeze(1)
