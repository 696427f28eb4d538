"""Harness that imports and runs the UNMODIFIED reference (TEST INFRASTRUCTURE ONLY).

Where the reference comes from: `oracle/_ref/` (the copy `tools/make_ref.py` makes, which
travels to the GPU box) or, in the build container, `/root/reference` itself.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this module; nothing under
`foundationstereo_b200/` does.

What the harness has to supply because the image lacks it, all
of it OUTSIDE the hot path and identical for the stock and the patched run:

* empty stand-ins for the third-party packages `Utils.py:10-18` imports at module top and the
  image lacks (`trimesh, open3d, imageio, transformations, omegaconf`): never called;
* `torch.hub.load('facebookresearch/dinov2', ...)` (`depth_anything/dpt.py:159`, a network call)
  redirected to the reference's own vendored `dinov2/` (source='local', pretrained=False);
* `timm.create_model('edgenext_small', pretrained=True)` (`core/extractor.py:327`; timm is not
  installed and there is no network): a DISCLOSED stand-in CNN with the contract the caller uses
  -- `.stem` (stride 4 -> 48 channels) and `.stages` (48/96/160/304 channels at strides
  4/8/16/32, `core/extractor.py:328-330,358-362`).  Weights are random-init anyway (BASELINE.json
  configs: "random init"), and parity is stock modules vs patched modules on the SAME object;
* `flash_attn_func` only takes 16-bit tensors, so the fp32 run of the reference that measures
  its own fp32-vs-autocast noise floor -- and any CPU dry run -- swaps it for the
  equivalent `scaled_dot_product_attention` (`core/submodule.py:224`, window (-1,-1) = full
  attention).  The stock-vs-patched comparison itself runs the real flash_attn on both sides.
"""
from __future__ import annotations

import contextlib
import importlib
import os
import sys
import types
import warnings
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
_STUBS = ("trimesh", "open3d", "imageio", "transformations", "omegaconf")
_state: Dict[str, object] = {}


def ref_root() -> Optional[str]:
    """oracle/_ref if it was made, else the build container's checkout, else None."""
    for p in (os.environ.get("FS_REFERENCE"), os.path.join(HERE, "_ref"), "/root/reference"):
        if p and os.path.isdir(os.path.join(p, "core")):
            return p
    return None


def available() -> bool:
    return ref_root() is not None


# ---------------------------------------------------------------------------------------------
# the timm stand-in (disclosed above)
# ---------------------------------------------------------------------------------------------
class _Block(nn.Module):
    def __init__(self, c: int):
        super().__init__()
        self.dw = nn.Conv2d(c, c, 7, padding=3, groups=c)
        self.norm = nn.GroupNorm(1, c)
        self.pw1 = nn.Conv2d(c, 4 * c, 1)
        self.pw2 = nn.Conv2d(4 * c, c, 1)

    def forward(self, x):
        return x + self.pw2(F.gelu(self.pw1(self.norm(self.dw(x)))))


class _EdgeNextStandIn(nn.Module):
    """`.stem`: (B,3,H,W) -> (B,48,H/4,W/4); `.stages[i]`: 48 -> 48 -> 96 -> 160 -> 304, stride 2 from i = 1."""

    def __init__(self, chans: Sequence[int] = (48, 96, 160, 304), depth: int = 2):
        super().__init__()
        self.stem = nn.Sequential(nn.Conv2d(3, chans[0], 4, stride=4), nn.GroupNorm(1, chans[0]))
        stages = []
        for i, c in enumerate(chans):
            layers: List[nn.Module] = []
            if i > 0:
                layers += [nn.GroupNorm(1, chans[i - 1]), nn.Conv2d(chans[i - 1], c, 2, stride=2)]
            layers += [_Block(c) for _ in range(depth)]
            stages.append(nn.Sequential(*layers))
        self.stages = nn.Sequential(*stages)


def _timm_shim() -> types.ModuleType:
    m = types.ModuleType("timm")
    m.__all__ = []
    m.__fsb_stand_in__ = True

    def create_model(name, pretrained=False, features_only=False, **kw):
        assert name == "edgenext_small", name
        return _EdgeNextStandIn()

    m.create_model = create_model
    return m


# ---------------------------------------------------------------------------------------------
# import
# ---------------------------------------------------------------------------------------------
def import_reference() -> types.SimpleNamespace:
    """Imports the reference's `core` package (once) and returns its hot-path symbols and modules."""
    if "ns" in _state:
        return _state["ns"]
    root = ref_root()
    if root is None:
        raise RuntimeError("the reference is not available: run tools/make_ref.py where /root/reference exists")
    os.environ.setdefault("XFORMERS_DISABLED", "1")
    warnings.filterwarnings("ignore")
    for name in _STUBS:
        try:
            importlib.import_module(name)
        except Exception:
            sys.modules[name] = types.ModuleType(name)
            sys.modules[name].__all__ = []
    try:
        importlib.import_module("timm")
    except Exception:
        sys.modules["timm"] = _timm_shim()
    for p in (os.path.join(root, "dinov2"), root):
        if p not in sys.path:
            sys.path.insert(0, p)
    # depth_anything/dpt.py:159 -> the vendored DINOv2, offline
    if not getattr(torch.hub.load, "__fsb_redirect__", False):
        real_load = torch.hub.load

        def load(repo_or_dir, model, *a, **kw):
            if repo_or_dir == "facebookresearch/dinov2":
                kw.pop("pretrained", None)
                kw.pop("source", None)
                return real_load(os.path.join(root, "dinov2"), model, *a, source="local", pretrained=False, **kw)
            return real_load(repo_or_dir, model, *a, **kw)

        load.__fsb_redirect__ = True
        torch.hub.load = load
    import logging
    level = logging.getLogger().level
    import core.submodule as submodule
    import core.geometry as geometry
    import core.update as update
    import core.utils.utils as utils
    logging.getLogger().setLevel(level)  # Utils.py:24-29 reconfigures root logging on import
    ns = types.SimpleNamespace(
        root=root, submodule=submodule, geometry=geometry, update=update, utils=utils,
        build_gwc_volume=submodule.build_gwc_volume, build_concat_volume=submodule.build_concat_volume,
        disparity_regression=submodule.disparity_regression, context_upsample=submodule.context_upsample,
        Combined_Geo_Encoding_Volume=geometry.Combined_Geo_Encoding_Volume,
        BasicSelectiveMultiUpdateBlock=update.BasicSelectiveMultiUpdateBlock,
        timm_stand_in=bool(getattr(sys.modules["timm"], "__fsb_stand_in__", False)))
    _state["ns"] = ns
    return ns


def import_model_module():
    """core.foundation_stereo (pulls in the extractor, DINOv2, DepthAnything)."""
    import_reference()
    import core.foundation_stereo as fs
    return fs


class Args(types.SimpleNamespace):
    """The `args` object FoundationStereo reads (attribute access + `.get`, like the SimpleNamespace
    subclass of scripts/train.py:45-64 and the OmegaConf of scripts/run_demo.py:112-117)."""

    def get(self, key, default=None):
        return getattr(self, key, default)

    def __contains__(self, key):
        return hasattr(self, key)

    def __getitem__(self, key):
        return getattr(self, key)


def model_args(vit_size: str = "vits", max_disp: int = 416, corr_levels: int = 2, corr_radius: int = 4,
               n_gru_layers: int = 3, mixed_precision: bool = True) -> Args:
    """configs/train/stereo_v1.json:2-16 with BASELINE.json's max_disp / vit_size."""
    return Args(max_disp=max_disp, hidden_dims=[128, 128, 128], n_gru_layers=n_gru_layers, n_downsample=2,
                corr_radius=corr_radius, corr_levels=corr_levels, mixed_precision=mixed_precision, vit_size=vit_size,
                low_memory=False, slow_fast_gru=False, corr_implementation="reg", shared_backbone=False)


def update_block_args(corr_levels: int = 2, corr_radius: int = 4, n_gru_layers: int = 3) -> Args:
    return Args(corr_levels=corr_levels, corr_radius=corr_radius, n_gru_layers=n_gru_layers)


def build_model(vit_size: str = "vits", max_disp: int = 416, corr_levels: int = 2, corr_radius: int = 4,
                mixed_precision: bool = True, seed: int = 0, delta_scale: float = 1.0, device="cpu") -> nn.Module:
    """Random-init `FoundationStereo` (core/foundation_stereo.py:127-181) in eval mode.
    `delta_scale` multiplies `update_block.disp_head.conv.4` (the layer that emits delta_disp,
    core/update.py:28): an untrained GRU is not contractive, 1.0 = untouched."""
    fs = import_model_module()
    torch.manual_seed(seed)
    model = fs.FoundationStereo(model_args(vit_size, max_disp, corr_levels, corr_radius, 3, mixed_precision))
    if delta_scale != 1.0:
        with torch.no_grad():
            model.update_block.disp_head.conv[4].weight.mul_(delta_scale)
            model.update_block.disp_head.conv[4].bias.mul_(delta_scale)
    return model.to(device).eval()


def make_pair(B: int, H: int, W: int, max_disp: int = 64, seed: int = 0) -> Tuple[torch.Tensor, torch.Tensor]:
    """(B,3,H,W) float images in 0..255: the left image is band-limited noise, the right image the
    left one warped by a smooth synthetic disparity in [2, max_disp] px so that matches exist
    (SURVEY.md 8(d))."""
    g = torch.Generator().manual_seed(seed)
    tex = 0
    for s in (4, 8, 16, 32):
        n = torch.rand(B, 3, -(-H // s) + 1, -(-W // s) + 1, generator=g)
        tex = tex + F.interpolate(n, size=(H, W), mode="bicubic", align_corners=True) / 4
    left = (tex.clamp(0, 1) * 255.0)
    d = torch.rand(B, 1, 4, 6, generator=g)
    d = 2.0 + (max_disp - 2.0) * F.interpolate(d, size=(H, W), mode="bicubic", align_corners=True).clamp(0, 1)
    xs = torch.arange(W, dtype=torch.float32).view(1, 1, 1, W).expand(B, 1, H, W)
    ys = torch.arange(H, dtype=torch.float32).view(1, 1, H, 1).expand(B, 1, H, W)
    # right(x) = left(x + d): sample the left image d px to the right
    gx = 2.0 * (xs + d) / (W - 1) - 1.0
    gy = 2.0 * ys / (H - 1) - 1.0
    grid = torch.cat([gx, gy], 1).permute(0, 2, 3, 1)
    right = F.grid_sample(left, grid, mode="bilinear", padding_mode="border", align_corners=True)
    return left.contiguous(), right.contiguous()


# ---------------------------------------------------------------------------------------------
# fp32 runs of the reference (noise floor): flash_attn_func only takes 16-bit tensors
# ---------------------------------------------------------------------------------------------
def _sdpa_attention(q, k, v, window_size=(-1, -1), **kw):
    """flash_attn_func(Q,K,V) on (B, L, heads, head_dim) with full attention == SDPA on (B, heads, L, d)."""
    assert tuple(window_size) == (-1, -1), "the reference only uses full attention (core/submodule.py:516-528)"
    o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2))
    return o.transpose(1, 2)


@contextlib.contextmanager
def sdpa_instead_of_flash_attn():
    ns = import_reference()
    saved = ns.submodule.flash_attn_func
    ns.submodule.flash_attn_func = _sdpa_attention
    try:
        yield
    finally:
        ns.submodule.flash_attn_func = saved


@contextlib.contextmanager
def strict_fp32():
    """cuDNN's allow_tf32 defaults to True: the reference's "fp32" convolutions on a B200 are TF32
    unless this is switched off."""
    a, b = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        yield
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = a, b


# ---------------------------------------------------------------------------------------------
# the reference's hot loop from recorded inputs (core/foundation_stereo.py:229-242, restated as a driver)
# ---------------------------------------------------------------------------------------------
class Recorder:
    """Records what the unchanged forward hands to the hot path: the constructor arguments of
    Combined_Geo_Encoding_Volume (:229) and the first update_block call's (net, inp, disp, att) (:240)."""

    def __init__(self, model):
        self.model = model
        self.geo_args = None
        self.first = None
        self.calls = 0

    def __enter__(self):
        fs = import_model_module()
        rec = self
        self._cls = fs.Combined_Geo_Encoding_Volume
        base = self._cls

        class Recording(base):
            def __init__(self, f1, f2, vol, num_levels=2, dx=None):
                rec.geo_args = (f1.detach().clone(), f2.detach().clone(), vol.detach().clone(), num_levels, dx)
                super().__init__(f1, f2, vol, num_levels=num_levels, dx=dx)

        fs.Combined_Geo_Encoding_Volume = Recording

        def pre(mod, args):
            if rec.calls == 0:
                net, inp, corr, disp, att = args
                rec.first = ([t.detach().clone() for t in net], [t.detach().clone() for t in inp],
                             disp.detach().clone(), [t.detach().clone() for t in att])
            rec.calls += 1

        self._hook = self.model.update_block.register_forward_pre_hook(pre)
        return self

    def __exit__(self, *exc):
        fs = import_model_module()
        fs.Combined_Geo_Encoding_Volume = self._cls
        self._hook.remove()


def reference_loop(update_block, geo_cls, geo_args, first, iters: int, autocast: bool,
                   trace: bool = False):
    """iters x [geo_fn(disp) -> update_block -> disp += delta] exactly as
    core/foundation_stereo.py:229-242 strings the reference's own modules together.
    Returns (disp (B,1,H,W) fp32, mask_feat, states) where `states` (with `trace`) holds the
    (net_list, disp) entering every iteration -- the teacher-forcing trajectory."""
    f1, f2, vol, L, dx = geo_args
    net0, inp, disp0, att = first
    dev = disp0.device
    geo_fn = geo_cls(f1.float(), f2.float(), vol.float(), num_levels=L, dx=dx)
    b, c, h, w = f1.shape
    coords = torch.arange(w, dtype=torch.float, device=dev).reshape(1, 1, w, 1).repeat(b, h, 1, 1)
    cast = (lambda t: t) if autocast else (lambda t: t.float())
    net_list = [cast(t.clone()) for t in net0]
    inp_list = [cast(t) for t in inp]
    att_list = [cast(t) for t in att]
    disp = disp0.float().clone()
    states = []
    mask = None
    for _ in range(iters):
        disp = disp.detach()
        if trace:
            states.append(([t.clone() for t in net_list], disp.clone()))
        geo_feat = geo_fn(disp, coords)
        with torch.autocast(dev.type, enabled=autocast):
            net_list, mask, delta = update_block(net_list, inp_list, geo_feat, disp, att_list)
        disp = disp + delta.float()
    return disp, mask, states


# ---------------------------------------------------------------------------------------------
# the reference's hot path on the CPU (bench.py's cpu_baseline / --impl reference legs)
# ---------------------------------------------------------------------------------------------
def make_reference_update_block(state: Dict[str, torch.Tensor], corr_levels: int = 2, corr_radius: int = 4,
                                n_gru_layers: int = 3):
    """The reference's BasicSelectiveMultiUpdateBlock (core/update.py:122-139) with `state` loaded strictly."""
    ns = import_reference()
    ub = ns.BasicSelectiveMultiUpdateBlock(update_block_args(corr_levels, corr_radius, n_gru_layers), 128, volume_dim=28)
    ub.load_state_dict(state, strict=True)
    return ub.eval()


def reference_hot_path_cpu(ub, x: Dict[str, object], max_disp: int, groups: int, corr_levels: int, corr_radius: int,
                           iters: int) -> Tuple[torch.Tensor, Dict[str, float]]:
    """One pair (or batch) through the reference's OWN functions, fp32 on the host cores:
    build_gwc_volume + build_concat_volume (core/submodule.py:399-427), torch.cat (:212 of the caller),
    Combined_Geo_Encoding_Volume.__init__ (core/geometry.py:18-40), iters x (__call__ :43-65 +
    BasicSelectiveMultiUpdateBlock.forward, core/update.py:141-159).  Returns the final 1/4-res disparity
    and the stage times in seconds."""
    import time
    ns = import_reference()
    D = max_disp // 4
    t = {}
    with torch.no_grad():
        t0 = time.perf_counter()
        gwc = ns.build_gwc_volume(x["fl"], x["fr"], D, groups)
        cat = ns.build_concat_volume(x["pl"], x["pr"], maxdisp=D)
        comb = torch.cat([gwc, cat], dim=1)
        t1 = time.perf_counter()
        dx = torch.linspace(-corr_radius, corr_radius, 2 * corr_radius + 1).reshape(1, 1, 2 * corr_radius + 1, 1)
        geo_fn = ns.Combined_Geo_Encoding_Volume(x["fl"].float(), x["fr"].float(), x["volume"].float(),
                                                 num_levels=corr_levels, dx=dx)
        t2 = time.perf_counter()
        b, c, h, w = x["fl"].shape
        coords = torch.arange(w, dtype=torch.float).reshape(1, 1, w, 1).repeat(b, h, 1, 1)
        disp = x["disp"].float()
        net = [n.clone() for n in x["net"]]
        for _ in range(iters):
            geo_feat = geo_fn(disp, coords)
            net, mask, delta = ub(net, list(x["inp"]), geo_feat, disp, list(x["att"]))
            disp = disp + delta.float()
        t3 = time.perf_counter()
    del comb
    t.update(volumes_s=t1 - t0, geo_init_s=t2 - t1, loop_s=t3 - t2, total_s=t3 - t0)
    return disp, t
