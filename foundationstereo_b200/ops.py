"""torch.library ops over the C ABI of libfsb200.so (kernels 1, 2, 2b, 3).

Each op checks device / dtype / contiguity, allocates its outputs with PyTorch
(so the caching allocator and stream semantics stay intact), and launches the
kernel on torch's current CUDA stream through ctypes.  The ops are CUDA-only and
inference-only: no autograd formula is registered, so differentiating through
them raises instead of silently returning zeros.
"""
from __future__ import annotations

from typing import List

import torch

from . import _lib
from ._lib import check, level_array, lib

_DTYPE = {torch.float32: _lib.FSB_F32, torch.float16: _lib.FSB_F16, torch.bfloat16: _lib.FSB_BF16}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts: torch.Tensor) -> None:
    """All tensors on one CUDA device."""
    for t in ts:
        if not t.is_cuda:
            raise _lib.FsbError("foundationstereo_b200 ops run on CUDA tensors only (no CPU path)")
        if t.device != ts[0].device:
            raise _lib.FsbError(f"tensors on different devices: {ts[0].device} and {t.device}")


def _call(t: torch.Tensor, name: str, *args) -> None:
    """One C-ABI launch on the current stream of ``t``'s device (which need not be the caller's current device)."""
    with torch.cuda.device(t.device):
        check(getattr(lib(), name)(*args, torch.cuda.current_stream().cuda_stream))


# ----------------------------------------------------------------------------------------
# kernel 1
# ----------------------------------------------------------------------------------------
@torch.library.custom_op("fsb200::gwc_volume", mutates_args=(), device_types="cuda")
def gwc_volume(fl: torch.Tensor, fr: torch.Tensor, maxdisp: int, num_groups: int) -> torch.Tensor:
    _need_cuda(fl, fr)
    assert fl.shape == fr.shape and fl.dtype == fr.dtype and fl.dim() == 4
    fl, fr = fl.contiguous(), fr.contiguous()
    B, C, H, W = fl.shape
    out = torch.empty((B, num_groups, maxdisp, H, W), dtype=fl.dtype, device=fl.device)
    _call(fl, "fsb_volume_gwc_concat", fl.data_ptr(), fr.data_ptr(), None, None, out.data_ptr(), None,
                                      B, C, 0, num_groups, H, W, maxdisp, _DTYPE[fl.dtype])
    return out


@gwc_volume.register_fake
def _(fl, fr, maxdisp, num_groups):
    B, C, H, W = fl.shape
    return fl.new_empty((B, num_groups, maxdisp, H, W))


@torch.library.custom_op("fsb200::concat_volume", mutates_args=(), device_types="cuda")
def concat_volume(pl: torch.Tensor, pr: torch.Tensor, maxdisp: int) -> torch.Tensor:
    _need_cuda(pl, pr)
    assert pl.shape == pr.shape and pl.dtype == pr.dtype and pl.dim() == 4
    pl, pr = pl.contiguous(), pr.contiguous()
    B, Cp, H, W = pl.shape
    out = torch.empty((B, 2 * Cp, maxdisp, H, W), dtype=pl.dtype, device=pl.device)
    _call(pl, "fsb_volume_gwc_concat", None, None, pl.data_ptr(), pr.data_ptr(), None, out.data_ptr(),
                                      B, 0, Cp, 0, H, W, maxdisp, _DTYPE[pl.dtype])
    return out


@concat_volume.register_fake
def _(pl, pr, maxdisp):
    B, Cp, H, W = pl.shape
    return pl.new_empty((B, 2 * Cp, maxdisp, H, W))


@torch.library.custom_op("fsb200::combined_volume", mutates_args=(), device_types="cuda")
def combined_volume(fl: torch.Tensor, fr: torch.Tensor, pl: torch.Tensor, pr: torch.Tensor,
                    maxdisp: int, num_groups: int) -> torch.Tensor:
    """gwc and concat volumes written by ONE launch into one (B, G+2Cp, D, H, W) tensor --
    what core/foundation_stereo.py:207-212 builds with two loops and a torch.cat."""
    _need_cuda(fl, fr, pl, pr)
    assert fl.dtype == fr.dtype == pl.dtype == pr.dtype
    fl, fr, pl, pr = fl.contiguous(), fr.contiguous(), pl.contiguous(), pr.contiguous()
    B, C, H, W = fl.shape
    Cp = pl.shape[1]
    out = torch.empty((B, num_groups + 2 * Cp, maxdisp, H, W), dtype=fl.dtype, device=fl.device)
    _call(fl, "fsb_volume_combined", fl.data_ptr(), fr.data_ptr(), pl.data_ptr(), pr.data_ptr(), out.data_ptr(),
                                    B, C, Cp, num_groups, H, W, maxdisp, _DTYPE[fl.dtype])
    return out


@combined_volume.register_fake
def _(fl, fr, pl, pr, maxdisp, num_groups):
    B, C, H, W = fl.shape
    return fl.new_empty((B, num_groups + 2 * pl.shape[1], maxdisp, H, W))


def volume_stem(fl: torch.Tensor, fr: torch.Tensor, pl: torch.Tensor, pr: torch.Tensor, weight: torch.Tensor,
                bias: torch.Tensor, maxdisp: int, num_groups: int) -> torch.Tensor:
    """corr_stem[0](cat([gwc volume, concat volume], 1)) in one launch (csrc/volstem.cu; core/foundation_stereo.py:207-213
    with the 1x1x1 Conv3d of :164-165): ``weight`` (Co, G + 2 Cp[, 1, 1, 1]) and ``bias`` (Co) are that convolution's
    parameters.  -> (B, Co, maxdisp, H, W) in the dtype of ``fl``.  Opt-in: the caller hands over part of corr_stem."""
    _need_cuda(fl, fr, pl, pr, weight, bias)
    assert fl.shape == fr.shape and pl.shape == pr.shape and fl.dtype == fr.dtype == pl.dtype == pr.dtype
    B, C, H, W = fl.shape
    Cp = pl.shape[1]
    w = weight.reshape(weight.shape[0], -1).float().contiguous()
    Co = w.shape[0]
    assert w.shape[1] == num_groups + 2 * Cp and bias.numel() == Co
    fl, fr, pl, pr, bias = fl.contiguous(), fr.contiguous(), pl.contiguous(), pr.contiguous(), bias.float().contiguous()
    out = fl.new_empty((B, Co, maxdisp, H, W))
    scratch = torch.empty((2, B, C, H, W), dtype=torch.float32, device=fl.device)   # the L2-normalised features
    wg_host = w[:, :num_groups].cpu().contiguous()   # (Co, G) on the host: passed to the kernel by value
    _call(fl, "fsb_volume_stem", fl.data_ptr(), fr.data_ptr(), pl.data_ptr(), pr.data_ptr(), w.data_ptr(), bias.data_ptr(),
          wg_host.data_ptr(), scratch.data_ptr(), out.data_ptr(), B, C, Cp, num_groups, Co, H, W, maxdisp, _DTYPE[fl.dtype])
    return out


# ----------------------------------------------------------------------------------------
# kernels 2 / 2b
# ----------------------------------------------------------------------------------------
@torch.library.custom_op("fsb200::corr_pyramid", mutates_args=(), device_types="cuda")
def corr_pyramid(f1: torch.Tensor, f2: torch.Tensor, num_levels: int, precision: int) -> List[torch.Tensor]:
    """levels l = 0..L-1 of the init correlation, each (B*H*W1, W2 >> l) fp32."""
    _need_cuda(f1, f2)
    f1 = f1.float().contiguous()
    f2 = f2.float().contiguous()
    B, C, H, W1 = f1.shape
    W2 = f2.shape[3]
    assert f2.shape[:3] == (B, C, H)
    levels = [torch.empty((B * H * W1, W2 >> l), dtype=torch.float32, device=f1.device) for l in range(num_levels)]
    _call(f1, "fsb_corr_pyramid", f1.data_ptr(), f2.data_ptr(), level_array(levels), B, C, H, W1, W2,
                                 num_levels, precision)
    return levels


@corr_pyramid.register_fake
def _(f1, f2, num_levels, precision):
    B, C, H, W1 = f1.shape
    W2 = f2.shape[3]
    return [f1.new_empty((B * H * W1, W2 >> l), dtype=torch.float32) for l in range(num_levels)]


@torch.library.custom_op("fsb200::geo_pyramid", mutates_args=(), device_types="cuda")
def geo_pyramid(volume: torch.Tensor, num_levels: int) -> List[torch.Tensor]:
    """levels of the geometry volume in the lookup layout (B*H*W, D >> l, Cg) fp32."""
    _need_cuda(volume)
    volume = volume.contiguous() if volume.dtype in _DTYPE else volume.float().contiguous()
    B, Cg, D, H, W = volume.shape
    levels = [torch.empty((B * H * W, D >> l, Cg), dtype=torch.float32, device=volume.device)
              for l in range(num_levels)]
    _call(volume, "fsb_geo_pyramid_in", volume.data_ptr(), _DTYPE[volume.dtype], level_array(levels), B, Cg, D, H, W,
          num_levels)
    return levels


@geo_pyramid.register_fake
def _(volume, num_levels):
    B, Cg, D, H, W = volume.shape
    return [volume.new_empty((B * H * W, D >> l, Cg), dtype=torch.float32) for l in range(num_levels)]


# ----------------------------------------------------------------------------------------
# kernel 3
# ----------------------------------------------------------------------------------------
@torch.library.custom_op("fsb200::geo_lookup", mutates_args=(), device_types="cuda")
def geo_lookup(geo_levels: List[torch.Tensor], corr_levels: List[torch.Tensor], disp: torch.Tensor,
               coords: torch.Tensor, radius: int) -> torch.Tensor:
    _need_cuda(disp, coords)
    L = len(geo_levels)
    assert len(corr_levels) == L
    B, _, H, W = disp.shape
    disp = disp.float().contiguous()
    coords = coords.float().contiguous()
    N, D, Cg = geo_levels[0].shape
    W2 = corr_levels[0].shape[1]
    assert N == B * H * W and coords.numel() == N
    K = 2 * radius + 1
    out = torch.empty((B, L * K * (Cg + 1), H, W), dtype=torch.float32, device=disp.device)
    _call(disp, "fsb_geo_lookup", level_array(geo_levels), level_array(corr_levels), disp.data_ptr(),
                               coords.data_ptr(), out.data_ptr(), B, Cg, D, H, W, W2, L, radius, 0)
    return out


@geo_lookup.register_fake
def _(geo_levels, corr_levels, disp, coords, radius):
    B, _, H, W = disp.shape
    Cg = geo_levels[0].shape[2]
    return disp.new_empty((B, len(geo_levels) * (2 * radius + 1) * (Cg + 1), H, W), dtype=torch.float32)


def geo_lookup_nhwc16(geo_levels: List[torch.Tensor], corr_levels: List[torch.Tensor], disp: torch.Tensor,
                      coords: torch.Tensor, radius: int, out: torch.Tensor) -> torch.Tensor:
    """Kernel 3 with the fused output layout: writes the samples as the 16-bit NHWC tensor
    ``out`` (B,H,W,Cpad), Cpad = L*K*(Cg+1) rounded up to 64 with zero pad channels -- the A
    operand of encoder.convc1.  ``out`` is caller-owned (a static buffer of the update-block
    plan), so the launch can be captured in a CUDA graph; not a torch.library op because it
    writes in place into that buffer."""
    _need_cuda(disp, coords, out)
    L = len(geo_levels)
    B, _, H, W = disp.shape
    N, D, Cg = geo_levels[0].shape
    W2 = corr_levels[0].shape[1]
    K = 2 * radius + 1
    cpad = (L * K * (Cg + 1) + 63) // 64 * 64
    assert out.shape == (B, H, W, cpad) and out.is_contiguous() and out.dtype in (torch.float16, torch.bfloat16)
    assert disp.dtype == torch.float32 and disp.is_contiguous() and coords.dtype == torch.float32 \
        and coords.is_contiguous() and coords.numel() == N == B * H * W
    layout = 1 if out.dtype == torch.float16 else 2
    _call(disp, "fsb_geo_lookup", level_array(geo_levels), level_array(corr_levels), disp.data_ptr(),
                               coords.data_ptr(), out.data_ptr(), B, Cg, D, H, W, W2, L, radius, layout)
    return out


def lookup_convc1(geo_levels: List[torch.Tensor], corr_levels: List[torch.Tensor], disp: torch.Tensor,
                  coords: torch.Tensor, radius: int, weight: torch.Tensor, bias: torch.Tensor,
                  out: torch.Tensor) -> torch.Tensor:
    """Kernels 3+4 fused (csrc/lookconv.cu): ``out`` (B,H,W,256) 16-bit NHWC = ReLU(convc1(lookup(disp))) without the
    lookup tensor ever reaching HBM (core/geometry.py:43-65 + core/update.py:56,63).  ``weight`` (256, Cpad) 16-bit =
    ``pack_convc1_weight(convc1.weight, L, K, Cg)``: level l's K*(Cg+1) columns at [l*Kl, ...), Kl = K*(Cg+1) rounded up
    to 64, zeros between; ``bias`` fp32."""
    _need_cuda(disp, coords, out, weight, bias)
    L = len(geo_levels)
    B, _, H, W = disp.shape
    N, D, Cg = geo_levels[0].shape
    W2 = corr_levels[0].shape[1]
    K = 2 * radius + 1
    cpad = L * ((K * (Cg + 1) + 63) // 64 * 64)
    cout = weight.shape[0]
    assert weight.shape == (cout, cpad) and weight.is_contiguous() and weight.dtype == out.dtype
    assert out.shape == (B, H, W, cout) and out.is_contiguous() and out.dtype in (torch.float16, torch.bfloat16)
    assert bias.dtype == torch.float32 and bias.numel() == cout
    assert disp.dtype == torch.float32 and disp.is_contiguous() and coords.dtype == torch.float32 \
        and coords.is_contiguous() and coords.numel() == N == B * H * W
    _call(disp, "fsb_lookup_convc1", level_array(geo_levels), level_array(corr_levels), disp.data_ptr(), coords.data_ptr(),
          weight.data_ptr(), bias.data_ptr(), out.data_ptr(), B, Cg, D, H, W, W2, L, radius, cout,
          1 if out.dtype == torch.float16 else 2)
    return out


def pack_convc1_weight(weight: torch.Tensor, L: int, K: int, Cg: int, dtype: torch.dtype) -> torch.Tensor:
    """convc1.weight (Cout, L*K*(Cg+1)[, 1, 1]) in the reference's channel order (core/geometry.py:60-65: per level the
    K*Cg geometry samples c*K + k, then the K correlation taps) -> the (Cout, L*Kl) 16-bit matrix fsb_lookup_convc1 reads:
    every level starts at a multiple of Kl = K*(Cg+1) rounded up to 64 columns, the columns in between are zero."""
    n = K * (Cg + 1)
    kl = (n + 63) // 64 * 64
    w = weight.reshape(weight.shape[0], -1)
    assert w.shape[1] == L * n, (tuple(weight.shape), L, K, Cg)
    out = torch.zeros(w.shape[0], L * kl, dtype=dtype, device=w.device)
    for l in range(L):
        out[:, l * kl:l * kl + n] = w[:, l * n:(l + 1) * n].to(dtype)
    return out.contiguous()


def corr_pyramid_into(f1: torch.Tensor, f2: torch.Tensor, levels: List[torch.Tensor], precision: int) -> None:
    """fsb_corr_pyramid into caller-owned level buffers (stable addresses for graph capture)."""
    _need_cuda(f1, f2)
    f1 = f1.float().contiguous()
    f2 = f2.float().contiguous()
    B, C, H, W1 = f1.shape
    W2 = f2.shape[3]
    for l, t in enumerate(levels):
        assert t.shape == (B * H * W1, W2 >> l) and t.dtype == torch.float32 and t.is_contiguous()
    _call(f1, "fsb_corr_pyramid", f1.data_ptr(), f2.data_ptr(), level_array(levels), B, C, H, W1, W2,
                                 len(levels), precision)


def geo_pyramid_into(volume: torch.Tensor, levels: List[torch.Tensor]) -> None:
    """fsb_geo_pyramid into caller-owned level buffers."""
    _need_cuda(volume)
    volume = volume.contiguous() if volume.dtype in _DTYPE else volume.float().contiguous()
    B, Cg, D, H, W = volume.shape
    for l, t in enumerate(levels):
        assert t.shape == (B * H * W, D >> l, Cg) and t.dtype == torch.float32 and t.is_contiguous()
    _call(volume, "fsb_geo_pyramid_in", volume.data_ptr(), _DTYPE[volume.dtype], level_array(levels), B, Cg, D, H, W,
          len(levels))


# ----------------------------------------------------------------------------------------
# the steps either side of the refinement loop (csrc/post.cu)
# ----------------------------------------------------------------------------------------
@torch.library.custom_op("fsb200::disparity_regression", mutates_args=(), device_types="cuda")
def disparity_regression(x: torch.Tensor, logits: bool) -> torch.Tensor:
    """(B,D,H,W) -> (B,1,H,W): sum_d x*d (core/submodule.py:431-435); ``logits`` fuses the
    softmax over D of core/foundation_stereo.py:218 (result fp32)."""
    _need_cuda(x)
    assert x.dim() == 4
    x = x.contiguous()
    B, D, H, W = x.shape
    out = torch.empty((B, 1, H, W), dtype=torch.float32 if logits else x.dtype, device=x.device)
    _call(x, "fsb_disparity_regression", x.data_ptr(), out.data_ptr(), B, D, H, W, int(logits), _DTYPE[x.dtype])
    return out


@disparity_regression.register_fake
def _(x, logits):
    B, D, H, W = x.shape
    return x.new_empty((B, 1, H, W), dtype=torch.float32 if logits else x.dtype)


@torch.library.custom_op("fsb200::context_upsample", mutates_args=(), device_types="cuda")
def context_upsample(disp_low: torch.Tensor, weights: torch.Tensor, logits: bool, scale: float) -> torch.Tensor:
    """(B,1,h,w) fp32 x (B,9,4h,4w) -> (B,4h,4w) fp32 (core/submodule.py:456-468); ``logits`` fuses
    the softmax over the 9 weights, ``scale`` multiplies the disparity (core/foundation_stereo.py:188-189)."""
    _need_cuda(disp_low, weights)
    B, c, h, w = disp_low.shape
    assert c == 1 and weights.shape == (B, 9, 4 * h, 4 * w), (disp_low.shape, weights.shape)
    disp_low = disp_low.float().contiguous()
    weights = weights.contiguous()
    out = torch.empty((B, 4 * h, 4 * w), dtype=torch.float32, device=disp_low.device)
    _call(disp_low, "fsb_context_upsample", disp_low.data_ptr(), weights.data_ptr(), out.data_ptr(), B, h, w, int(logits),
                                     float(scale), _DTYPE[weights.dtype])
    return out


@context_upsample.register_fake
def _(disp_low, weights, logits, scale):
    B, _, h, w = disp_low.shape
    return disp_low.new_empty((B, 4 * h, 4 * w), dtype=torch.float32)


# ----------------------------------------------------------------------------------------
# banded init correlation (opt-in; include/fsb200.h "BANDED form")
# ----------------------------------------------------------------------------------------
def corr_band_geometry(d_band: int, radius: int, num_levels: int):
    """(bw0, off, align) of the band for lookups with disparities in [0, d_band] (fsb_corr_band_geometry)."""
    import ctypes
    bw0, off, align = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    check(lib().fsb_corr_band_geometry(int(d_band), int(radius), int(num_levels), ctypes.byref(bw0), ctypes.byref(off),
                                       ctypes.byref(align)))
    return bw0.value, off.value, align.value


def corr_pyramid_band(f1: torch.Tensor, f2: torch.Tensor, num_levels: int, precision: int, d_band: int, radius: int,
                      levels: List[torch.Tensor] | None = None) -> List[torch.Tensor]:
    """Banded correlation pyramid: level l is (B*H*W, bw0 >> l) fp32 (see include/fsb200.h).  ``levels``: optional
    caller-owned buffers of those shapes."""
    _need_cuda(f1, f2)
    f1 = f1.float().contiguous()
    f2 = f2.float().contiguous()
    B, C, H, W = f1.shape
    assert f2.shape == f1.shape, "the band form needs left and right features of one shape"
    bw0, _, _ = corr_band_geometry(d_band, radius, num_levels)
    if levels is None:
        levels = [torch.empty((B * H * W, bw0 >> l), dtype=torch.float32, device=f1.device) for l in range(num_levels)]
    for l, t in enumerate(levels):
        assert t.shape == (B * H * W, bw0 >> l) and t.dtype == torch.float32 and t.is_contiguous()
    _call(f1, "fsb_corr_pyramid_band", f1.data_ptr(), f2.data_ptr(), level_array(levels), B, C, H, W, num_levels, precision,
          int(d_band), int(radius))
    return levels


def geo_lookup_band(geo_levels: List[torch.Tensor], corr_levels: List[torch.Tensor], disp: torch.Tensor,
                    coords: torch.Tensor, radius: int, d_band: int, out16: torch.Tensor | None = None) -> torch.Tensor:
    """Kernel 3 over a banded correlation pyramid: fp32 (B, L*K*(Cg+1), H, W), or into the 16-bit NHWC buffer ``out16``."""
    _need_cuda(disp, coords)
    L = len(geo_levels)
    B, _, H, W = disp.shape
    disp = disp.float().contiguous()
    coords = coords.float().contiguous()
    N, D, Cg = geo_levels[0].shape
    assert N == B * H * W and coords.numel() == N and corr_levels[0].shape[0] == N
    K = 2 * radius + 1
    if out16 is None:
        out = torch.empty((B, L * K * (Cg + 1), H, W), dtype=torch.float32, device=disp.device)
        layout = 0
    else:
        cpad = (L * K * (Cg + 1) + 63) // 64 * 64
        assert out16.shape == (B, H, W, cpad) and out16.is_contiguous() and out16.dtype in (torch.float16, torch.bfloat16)
        out, layout = out16, (1 if out16.dtype == torch.float16 else 2)
    _call(disp, "fsb_geo_lookup_band", level_array(geo_levels), level_array(corr_levels), disp.data_ptr(), coords.data_ptr(),
          out.data_ptr(), B, Cg, D, H, W, L, radius, layout, int(d_band))
    return out
