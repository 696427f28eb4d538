"""Installs the B200 path into an UNCHANGED copy of the reference.

The reference has no plugin API; its seam is Python name binding:
  * core/foundation_stereo.py:19 star-imports build_gwc_volume / build_concat_volume,
    :18 imports Combined_Geo_Encoding_Volume by name -> the names live in that module's
    globals and are looked up at call time (:207, :210, :229);
  * the update block is the attribute ``model.update_block`` (:137), called at :240; its
    parameters must stay where they are so checkpoints load (scripts/run_demo.py:125);
  * the steps either side of the loop, disparity_regression (:220) and context_upsample (:189),
    come in through the same star import.
"""
from __future__ import annotations

import sys
import types
from typing import Optional

from .geometry import Combined_Geo_Encoding_Volume
from .submodule import (build_concat_volume, build_gwc_volume, context_upsample, disparity_regression,
                        softmax_context_upsample)
from .update import SelectiveUpdateBlock

_SAVED = {}


def install(model=None, modules: Optional[list] = None, fused_lookup: bool = True,
            fused_upsample: bool = False, corr_band: Optional[int] = None, fused_stem: bool = False) -> None:
    """Rebind the three names in every loaded reference module that holds them and, if a
    FoundationStereo ``model`` is given, route ``model.update_block.forward`` through
    SelectiveUpdateBlock built from the module's own parameters.

    With a model and ``fused_lookup`` (default) ``geo_fn(disp, coords)`` writes its samples
    straight into the update block's 16-bit NHWC input buffer and returns a channels-last
    view of it -- the unchanged caller (core/foundation_stereo.py:238-240) only hands that
    tensor on to ``update_block``.  ``fused_lookup=False`` returns the reference's fp32
    (B, L*K*(Cg+1), H, W) tensor.

    ``fused_upsample`` (opt-in, needs a model) swaps ``model.upsample_disp``
    (core/foundation_stereo.py:183-191) for a method that runs the two spx networks as before and
    then ONE kernel for softmax over the 9 weights + ``disp*4`` + convex upsampling.

    ``fused_stem`` (opt-in, needs a model) folds the first layer of ``model.corr_stem`` -- the 1x1x1
    Conv3d(32 -> 28) of core/foundation_stereo.py:164-165 -- into the volume kernel (csrc/volstem.cu): the unchanged
    caller's ``build_gwc_volume`` call (:207) only records the features and returns a zero-channel tensor, its
    ``build_concat_volume`` call (:210) returns the convolution's RESULT, ``torch.cat`` (:212) passes it through and
    ``corr_stem[0].forward`` becomes the identity (the module and its ``state_dict`` keys stay).  The 32-channel volume is
    never written, concatenated or read back."""
    geo_cls = Combined_Geo_Encoding_Volume
    if corr_band:
        # opt-in: banded init correlation for disparities in [0, corr_band] (use model.args.max_disp // 4); lookups
        # outside that range read zeros where the reference reads real correlations (include/fsb200.h)
        class BandedGeoEncodingVolume(Combined_Geo_Encoding_Volume):
            def __init__(self, init_fmap1, init_fmap2, geo_volume, num_levels=2, dx=None, **kw):
                super().__init__(init_fmap1, init_fmap2, geo_volume, num_levels=num_levels, dx=dx, corr_band=corr_band, **kw)
        geo_cls = BandedGeoEncodingVolume
    geo_base = geo_cls
    if model is not None:
        engine_of = install_update_block(model.update_block)
        if fused_lookup:
            class FusedGeoEncodingVolume(geo_base):
                def __init__(self, init_fmap1, init_fmap2, geo_volume, num_levels=2, dx=None):
                    super().__init__(init_fmap1, init_fmap2, geo_volume, num_levels=num_levels, dx=dx,
                                     nhwc16_out=lambda B, H, W: engine_of().corr_buffer(B, H, W))
            FusedGeoEncodingVolume.__name__ = "Combined_Geo_Encoding_Volume"
            geo_cls = FusedGeoEncodingVolume
        if fused_upsample:
            install_upsample(model)
    gwc_fn, cat_fn = build_gwc_volume, build_concat_volume
    if fused_stem:
        if model is None:
            raise ValueError("fused_stem needs the model (it reads corr_stem[0]'s parameters)")
        gwc_fn, cat_fn = install_stem(model)
    names = {
        "build_gwc_volume": gwc_fn,
        "build_concat_volume": cat_fn,
        "Combined_Geo_Encoding_Volume": geo_cls,
        "disparity_regression": disparity_regression,
        "context_upsample": context_upsample,
    }
    if modules is None:
        modules = [m for n, m in list(sys.modules.items())
                   if m is not None and n in ("core.foundation_stereo", "core.submodule", "core.geometry")]
    for mod in modules:
        for name, fn in names.items():
            if hasattr(mod, name):
                _SAVED.setdefault((mod.__name__, name), getattr(mod, name))
                setattr(mod, name, fn)


def install_update_block(update_block):
    """Keep the nn.Module (and its state_dict keys) in place, replace only ``forward``.
    The engine re-reads the parameters when their version counters change.  Returns a
    callable that yields the current engine (built on first use)."""
    state = {"engine": None, "versions": None}

    def engine():
        # in-place updates bump `_version`; `.to()` / `.half()` / `.cuda()` swap `param.data` without bumping it, so the
        # storage address, device and dtype are part of the key as well
        versions = tuple((p._version, p.data_ptr(), p.device, p.dtype) for p in update_block.parameters())
        if state["engine"] is None or versions != state["versions"]:
            dev = next(update_block.parameters()).device
            state["engine"] = None  # drop the old engine (packed weights, plans, captured graphs) first
            state["engine"] = SelectiveUpdateBlock(update_block.state_dict(), dev)
            state["versions"] = versions
        return state["engine"]

    def fast_forward(self, net, inp, corr, disp, att):
        import torch
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            from ._lib import FsbError
            raise FsbError("the B200 update block is inference-only: with gradients enabled its weights would silently "
                           "stop training.  Run under torch.no_grad() / torch.inference_mode() (scripts/run_demo.py:102 "
                           "does), or fsb.uninstall(model) before training")
        return engine()(net, inp, corr, disp, att)

    _SAVED.setdefault(("update_block", id(update_block)), update_block.forward)
    update_block.forward = types.MethodType(fast_forward, update_block)
    return engine


def install_stem(model):
    """-> (build_gwc_volume, build_concat_volume) stand-ins that together return corr_stem[0](cat([gwc, concat], 1)) from
    ONE launch (ops.volume_stem), and make ``model.corr_stem[0].forward`` the identity."""
    from . import ops
    conv0 = model.corr_stem[0]
    pending = {}

    def gwc_deferred(refimg_fea, targetimg_fea, maxdisp, num_groups, stride=1):
        pending["f"] = (refimg_fea, targetimg_fea, num_groups)
        B, _, H, W = refimg_fea.shape
        return refimg_fea.new_empty((B, 0, maxdisp, H, W))   # torch.cat (core/foundation_stereo.py:212) skips it

    def concat_with_stem(refimg_fea, targetimg_fea, maxdisp):
        fl, fr, groups = pending.pop("f")
        dt = fl.dtype
        return ops.volume_stem(fl, fr, refimg_fea.to(dt), targetimg_fea.to(dt), conv0.weight, conv0.bias, maxdisp, groups)

    _SAVED.setdefault(("corr_stem0", id(model)), conv0.forward)
    conv0.forward = types.MethodType(lambda self, x: x, conv0)
    return gwc_deferred, concat_with_stem


def install_upsample(model) -> None:
    """``model.upsample_disp`` (core/foundation_stereo.py:183-191) with its tail fused: the spx
    networks stay the model's own modules, softmax + ``disp*4.`` + context_upsample are one launch."""
    import torch

    def fast_upsample_disp(self, disp, mask_feat_4, stem_2x):
        with torch.autocast("cuda", enabled=bool(self.args.mixed_precision)):
            xspx = self.spx_2_gru(mask_feat_4, stem_2x)
            spx_pred = self.spx_gru(xspx)
        return softmax_context_upsample(disp.float(), spx_pred, 4.0)

    _SAVED.setdefault(("upsample_disp", id(model)), model.upsample_disp)
    model.upsample_disp = types.MethodType(fast_upsample_disp, model)


def uninstall(model=None) -> None:
    for (modname, name), fn in list(_SAVED.items()):
        if modname in ("update_block", "upsample_disp", "corr_stem0"):
            continue
        mod = sys.modules.get(modname)
        if mod is not None:
            setattr(mod, name, fn)
    if model is not None:
        ub = getattr(model, "update_block", None)
        key = ("update_block", id(ub))
        if ub is not None and key in _SAVED:
            ub.forward = _SAVED.pop(key)
        key = ("upsample_disp", id(model))
        if key in _SAVED:
            model.upsample_disp = _SAVED.pop(key)
        key = ("corr_stem0", id(model))
        if key in _SAVED:
            model.corr_stem[0].forward = _SAVED.pop(key)
