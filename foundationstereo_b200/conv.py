"""Host side of kernel 4 (csrc/conv.cu): weight packing for the tcgen05 implicit-GEMM
convolution and thin launch wrappers for it and for the NHWC helpers (csrc/nhwc.cu).

Activations are NHWC 16-bit tensors whose channel count is a multiple of 64; a
convolution reads the channel-concatenation of up to four such tensors ("sources"), which
is how the torch.cat calls of core/update.py:67,69,92,94,113,115,151 disappear.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import ConvArgs, check, lib

Tensor = torch.Tensor
KBLK = 64  # channels per K step (128 bytes of 16-bit elements)

_DT = {torch.float16: _lib.FSB_F16, torch.bfloat16: _lib.FSB_BF16}


def pad64(c: int) -> int:
    return (c + KBLK - 1) // KBLK * KBLK


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def pack_segment(w: Tensor, src_channels: Sequence[Tuple[int, int]], layout: int = 0) -> Tensor:
    """(Cout, Cin, k, k) -> (Cout, k*k*sum(Cpad)) in the K order of the kernel.
    layout 0 (FSB_CONV_LAYOUT_TAP): tap (ky, kx) row-major, then source, then channel.
    layout 1 (FSB_CONV_LAYOUT_HALO): source, then 64-channel chunk, then kx, then ky, then channel.
    ``src_channels`` lists (real, padded) channels of each source in concatenation order;
    pad channels get zero weights."""
    if w.dim() == 2:  # nn.Linear == 1x1 convolution
        w = w[:, :, None, None]
    cout, cin, kh, kw = w.shape
    assert sum(c for c, _ in src_channels) == cin, (w.shape, src_channels)
    wt = w.permute(0, 2, 3, 1)  # (Cout, kh, kw, Cin)
    parts, o = [], 0
    for real, padded in src_channels:
        part = wt[..., o:o + real]
        if padded > real:
            part = torch.nn.functional.pad(part, (0, padded - real))
        if layout == _lib.FSB_CONV_LAYOUT_HALO:
            assert padded % KBLK == 0
            part = part.reshape(cout, kh, kw, padded // KBLK, KBLK).permute(0, 3, 2, 1, 4).reshape(cout, -1)
        parts.append(part)
        o += real
    if layout == _lib.FSB_CONV_LAYOUT_HALO:
        return torch.cat(parts, 1)
    return torch.cat(parts, -1).reshape(cout, -1)


def pack_rows(mats: Sequence[Tensor], rows: int) -> Tensor:
    """Concatenate per-segment (Cout, K_s) matrices along K and zero-pad to ``rows`` rows."""
    m = torch.cat(list(mats), 1)
    if m.shape[0] < rows:
        m = torch.nn.functional.pad(m, (0, 0, 0, rows - m.shape[0]))
    return m.contiguous()


def interleave_zr(z: Tensor, r: Tensor, block_n: int) -> Tensor:
    """Row order of a GATE_ZR launch: tile t = [z rows of hidden channels t*bn/2 .. | r rows of
    the same channels], so one CTA owns both gates of its channels (csrc/conv.cu epilogue)."""
    half = block_n // 2
    hidden = z.shape[0]
    assert hidden % half == 0
    out = []
    for t in range(hidden // half):
        out += [z[t * half:(t + 1) * half], r[t * half:(t + 1) * half]]
    return torch.cat(out, 0)


class ConvLaunch:
    """One prepared fsb_conv2d_nhwc call: the argument struct is filled once (buffers are
    static), so a launch is a single FFI call and can be captured in a CUDA graph."""

    def __init__(self, *, srcs: Sequence[Tensor], segs: Sequence[Tuple[int, Sequence[int]]], weight: Tensor,
                 bias: Tensor, cout: int, block_n: int, epilogue: int, out: Optional[Tensor] = None, ch_off: int = 0,
                 out_f32: Optional[Tensor] = None, out_scale: float = 1.0, res: Optional[Tensor] = None,
                 gamma: Optional[Tensor] = None, h32: Optional[Tensor] = None, att: Optional[Tensor] = None,
                 z_l: Optional[Tensor] = None, z_s: Optional[Tensor] = None, rh_l: Optional[Tensor] = None,
                 rh_s: Optional[Tensor] = None, h32_out: Optional[Tensor] = None, h16_out: Optional[Tensor] = None,
                 hidden: int = 0, stages: int = 0, layout: int = 0, name: str = ""):
        a = ConvArgs()
        B, H, W, _ = srcs[0].shape
        self.dtype = _DT[srcs[0].dtype]
        for i, s in enumerate(srcs):
            assert s.is_cuda and s.is_contiguous() and s.shape[:3] == (B, H, W) and s.shape[3] % KBLK == 0, \
                (name, i, tuple(s.shape))
            a.src[i].ptr = s.data_ptr()
            a.src[i].C = s.shape[3]
        a.nsrc = len(srcs)
        a.B, a.H, a.W = B, H, W
        ktot = 0
        for j, (ksize, idx) in enumerate(segs):
            a.seg[j].ksize = ksize
            a.seg[j].nsrc = len(idx)
            for k, si in enumerate(idx):
                a.seg[j].src[k] = si
                ktot += ksize * ksize * srcs[si].shape[3]
        a.nseg = len(segs)
        rows = weight.shape[0]
        assert weight.dtype == srcs[0].dtype and weight.is_contiguous() and weight.shape == (rows, ktot), \
            (name, tuple(weight.shape), rows, ktot)
        assert bias.dtype == torch.float32 and bias.numel() == len(segs) * rows, (name, bias.shape, rows)
        a.weight, a.bias = weight.data_ptr(), bias.data_ptr()
        a.rows, a.Cout, a.block_n, a.stages, a.epilogue = rows, cout, block_n, stages, epilogue
        a.layout = layout
        ptr = lambda t: t.data_ptr() if t is not None else None
        if out is not None:
            a.out, a.ld_out, a.ch_off = out.data_ptr(), out.shape[3], ch_off
        a.out_f32, a.out_scale = ptr(out_f32), out_scale
        if res is not None:
            a.res, a.ld_res = res.data_ptr(), res.shape[3]
        a.gamma, a.h32, a.att = ptr(gamma), ptr(h32), ptr(att)
        a.z_l, a.z_s, a.rh_l, a.rh_s = ptr(z_l), ptr(z_s), ptr(rh_l), ptr(rh_s)
        a.h32_out, a.h16_out, a.hidden = ptr(h32_out), ptr(h16_out), hidden
        self.args = a
        self.name = name
        self.flops = 2.0 * B * H * W * rows * ktot
        # keep every tensor alive as long as the launch exists
        self._keep = [srcs, weight, bias, out, out_f32, res, gamma, h32, att, z_l, z_s, rh_l, rh_s, h32_out, h16_out]
        self._fn = lib().fsb_conv2d_nhwc
        self._ref = ctypes.byref(a)

    def __call__(self) -> None:
        check(self._fn(self._ref, self.dtype, _stream()))


# ---------------------------------------------------------------------------------------
# helpers (csrc/nhwc.cu)
# ---------------------------------------------------------------------------------------
def nchw_to_nhwc(src: Tensor, dst16: Optional[Tensor], dst32: Optional[Tensor]) -> None:
    """(B,C,H,W) fp32 contiguous -> dst16 (B,H,W,Cpad) 16-bit and/or dst32 (B,H,W,C) fp32."""
    B, C, H, W = src.shape
    assert src.dtype == torch.float32 and src.is_contiguous()
    cpad = dst16.shape[3] if dst16 is not None else C
    dt = _DT[dst16.dtype] if dst16 is not None else _lib.FSB_F16
    check(lib().fsb_nchw_to_nhwc(src.data_ptr(), dst16.data_ptr() if dst16 is not None else None,
                                 dst32.data_ptr() if dst32 is not None else None, B, C, H, W, cpad, dt, _stream()))


def nhwc_to_nchw(src: Tensor, dst: Tensor) -> None:
    B, H, W, C = src.shape
    check(lib().fsb_nhwc_to_nchw(src.data_ptr(), dst.data_ptr(), B, C, H, W, _stream()))


def pool2x(src: Tensor, dst16: Tensor) -> None:
    """pool2x (core/update.py:72-73) of an NHWC tensor; ``src`` fp32 or the 16-bit type of ``dst16``."""
    B, H, W, C = src.shape
    assert dst16.shape == (B, (H + 1) // 2, (W + 1) // 2, C)
    if src.dtype == torch.float32:
        check(lib().fsb_pool2x_nhwc(src.data_ptr(), dst16.data_ptr(), B, C, H, W, _DT[dst16.dtype], _stream()))
    else:
        assert src.dtype == dst16.dtype
        check(lib().fsb_pool2x_nhwc16(src.data_ptr(), dst16.data_ptr(), B, C, H, W, _DT[dst16.dtype], _stream()))


def resize(src: Tensor, dst16: Tensor) -> None:
    """interp (core/update.py:78-80) of an NHWC tensor; ``src`` fp32 or the 16-bit type of ``dst16``."""
    B, Hs, Ws, C = src.shape
    _, Hd, Wd, _ = dst16.shape
    if src.dtype == torch.float32:
        check(lib().fsb_resize_nhwc(src.data_ptr(), dst16.data_ptr(), B, C, Hs, Ws, Hd, Wd, _DT[dst16.dtype], _stream()))
    else:
        assert src.dtype == dst16.dtype
        check(lib().fsb_resize_nhwc16(src.data_ptr(), dst16.data_ptr(), B, C, Hs, Ws, Hd, Wd, _DT[dst16.dtype], _stream()))


def conv7x7_c1(disp: Tensor, weight: Tensor, bias: Tensor, out: Tensor, cout: int, disp_out: Optional[Tensor],
               disp_ch: int) -> None:
    B, _, H, W = disp.shape
    check(lib().fsb_conv7x7_c1(disp.data_ptr(), weight.data_ptr(), bias.data_ptr(), out.data_ptr(), out.shape[3], cout,
                               disp_out.data_ptr() if disp_out is not None else None,
                               disp_out.shape[3] if disp_out is not None else 0, disp_ch, B, H, W, _DT[out.dtype],
                               _stream()))


def dwconv7x7(x: Tensor, weight: Tensor, bias: Tensor, out: Tensor) -> None:
    B, H, W, C = x.shape
    check(lib().fsb_dwconv7x7_nhwc(x.data_ptr(), weight.data_ptr(), bias.data_ptr(), out.data_ptr(), B, C, H, W,
                                   _DT[x.dtype], _stream()))


def edgenext_mlp(x: Tensor, w1: Tensor, b1: Tensor, w2: Tensor, b2: Tensor, gamma: Tensor, res: Tensor,
                 out: Tensor) -> None:
    """out = res + gamma * (GELU(x @ w1^T + b1) @ w2^T + b2) on 16-bit NHWC tensors (csrc/mlp.cu)."""
    C, F = w1.shape[1], w1.shape[0]
    assert x.is_contiguous() and res.is_contiguous() and out.is_contiguous() and x.shape == res.shape == out.shape
    assert w1.is_contiguous() and w2.is_contiguous() and w2.shape == (C, F) and x.shape[-1] == C
    assert x.dtype == w1.dtype == w2.dtype == res.dtype == out.dtype and b1.dtype == b2.dtype == gamma.dtype == torch.float32
    check(lib().fsb_edgenext_mlp(x.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(), gamma.data_ptr(),
                                 res.data_ptr(), out.data_ptr(), x.numel() // C, C, F, _DT[x.dtype], _stream()))


def conv3x3_cout1(x: Tensor, weight: Tensor, bias: Tensor, delta: Tensor, disp_accum: Optional[Tensor]) -> None:
    """delta = conv3x3(x; C -> 1) + bias on CUDA cores (csrc/nhwc.cu); with ``disp_accum`` also disp_accum += delta.
    x 16-bit NHWC (B,H,W,C); weight fp32 (9, C); delta / disp_accum fp32 (B,1,H,W)."""
    B, H, W, C = x.shape
    assert weight.dtype == torch.float32 and weight.shape == (9, C) and weight.is_contiguous()
    assert delta.dtype == torch.float32 and delta.numel() == B * H * W
    assert disp_accum is None or (disp_accum.dtype == torch.float32 and disp_accum.numel() == B * H * W)
    check(lib().fsb_conv3x3_cout1(x.data_ptr(), weight.data_ptr(), bias.data_ptr(), delta.data_ptr(),
                                  disp_accum.data_ptr() if disp_accum is not None else None, B, C, H, W, _DT[x.dtype],
                                  _stream()))


def edgenext_block(x: Tensor, dw_w: Tensor, dw_b: Tensor, w1: Tensor, b1: Tensor, w2: Tensor, b2: Tensor, gamma: Tensor,
                   out: Tensor) -> None:
    """out = x + gamma * (GELU(dwconv7x7(x) @ w1^T + b1) @ w2^T + b2): one EdgeNeXt block in one launch (csrc/edgenext.cu)."""
    B, H, W, C = x.shape
    F = w1.shape[0]
    assert x.is_contiguous() and out.is_contiguous() and x.shape == out.shape and x.data_ptr() != out.data_ptr()
    assert w1.is_contiguous() and w2.is_contiguous() and w1.shape == (F, C) and w2.shape == (C, F)
    assert x.dtype == w1.dtype == w2.dtype == out.dtype and dw_w.dtype == torch.float32 and dw_w.shape == (C, 49)
    check(lib().fsb_edgenext_block(x.data_ptr(), dw_w.data_ptr(), dw_b.data_ptr(), w1.data_ptr(), b1.data_ptr(),
                                   w2.data_ptr(), b2.data_ptr(), gamma.data_ptr(), out.data_ptr(), B, H, W, C, F,
                                   _DT[x.dtype], _stream()))


def edgenext_block_noalloc(x: Tensor, dw_w: Tensor, dw_b: Tensor, w1: Tensor, b1: Tensor, w2: Tensor, b2: Tensor,
                           gamma: Tensor, out: Tensor) -> None:
    """``edgenext_block`` for callers that cannot allocate: the depthwise output lives in a static device array of the
    library (csrc/noalloc.cu; at most ``lib().fsb_noalloc_max_pixels()`` pixels per call, calls that overlap on the GPU
    race on it).  No counterpart in the reference."""
    B, H, W, C = x.shape
    F = w1.shape[0]
    assert x.is_contiguous() and out.is_contiguous() and x.shape == out.shape and x.data_ptr() != out.data_ptr()
    assert w1.is_contiguous() and w2.is_contiguous() and w1.shape == (F, C) and w2.shape == (C, F)
    assert x.dtype == w1.dtype == w2.dtype == out.dtype and dw_w.dtype == torch.float32 and dw_w.shape == (C, 49)
    check(lib().fsb_edgenext_block_noalloc(x.data_ptr(), dw_w.data_ptr(), dw_b.data_ptr(), w1.data_ptr(), b1.data_ptr(),
                                           w2.data_ptr(), b2.data_ptr(), gamma.data_ptr(), out.data_ptr(), B, H, W, C, F,
                                           _DT[x.dtype], _stream()))


def axpy(y: Tensor, x: Tensor) -> None:
    assert y.dtype == x.dtype == torch.float32 and y.numel() == x.numel()
    check(lib().fsb_axpy_f32(y.data_ptr(), x.data_ptr(), y.numel(), _stream()))
