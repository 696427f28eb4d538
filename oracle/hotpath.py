"""Plain fp32 PyTorch-on-CPU restatement of the reference hot path.

Every function cites the reference lines it restates (paths relative to the
reference checkout).  The code is written from the algorithm, not transcribed:
loops are vectorised where that does not change the arithmetic, the sampler is
a direct 1-D linear interpolation instead of ``grid_sample``, and the update
block is functional over a flat ``{name: tensor}`` parameter dict whose keys
are the reference module's ``state_dict`` keys.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
Params = Dict[str, Tensor]


# --------------------------------------------------------------------------
# volumes  (core/submodule.py:388-427)
# --------------------------------------------------------------------------
def l2_normalize(x: Tensor, dim: int, eps: float = 1e-12) -> Tensor:
    """x / max(||x||_2, eps) along ``dim`` -- what F.normalize does
    (used at core/submodule.py:395 and core/geometry.py:75)."""
    n = torch.linalg.vector_norm(x, ord=2, dim=dim, keepdim=True)
    return x / n.clamp_min(eps)


def gwc_volume(left: Tensor, right: Tensor, maxdisp: int, num_groups: int) -> Tensor:
    """Group-wise correlation volume (core/submodule.py:388-412).

    V[b,g,d,h,w] = sum_c Lhat[b,g,c,h,w] * Rhat[b,g,c,h,w-d] for w >= d, else 0,
    with Lhat/Rhat L2-normalised over the C/G channels of each group, math in
    fp32, result stored in the dtype of ``left``.

    The reference re-normalises inside its per-disparity loop; the norm of a
    pixel's group vector does not depend on d, so normalising once gives the
    same values (bit-exact on CPU).
    """
    B, C, H, W = left.shape
    assert C % num_groups == 0, (C, num_groups)
    cpg = C // num_groups
    lh = l2_normalize(left.float().reshape(B, num_groups, cpg, H, W), 2)
    rh = l2_normalize(right.float().reshape(B, num_groups, cpg, H, W), 2)
    vol = left.new_zeros((B, num_groups, maxdisp, H, W))
    for d in range(min(maxdisp, W)):
        vol[:, :, d, :, d:] = (lh[..., d:] * rh[..., : W - d]).sum(2).to(vol.dtype)
    return vol


def concat_volume(left: Tensor, right: Tensor, maxdisp: int) -> Tensor:
    """Concatenation volume (core/submodule.py:416-427).

    First C channels: left features at every disparity and every w (also
    w < d).  Last C channels: right features shifted by d, zero for w < d.
    """
    B, C, H, W = left.shape
    vol = left.new_zeros((B, 2 * C, maxdisp, H, W))
    vol[:, :C] = left[:, :, None]
    for d in range(min(maxdisp, W)):
        vol[:, C:, d, :, d:] = right[..., : W - d]
    return vol


def volume_stem(left: Tensor, right: Tensor, proj_left: Tensor, proj_right: Tensor, weight: Tensor, bias: Tensor,
                maxdisp: int, num_groups: int) -> Tensor:
    """What the caller computes up to the first layer of corr_stem (core/foundation_stereo.py:207-213 with the 1x1x1
    Conv3d(32 -> 28) of :164-165): corr_stem[0](cat([gwc volume, concat volume], 1)).  ``weight`` (Co, G + 2 Cp[,1,1,1]),
    ``bias`` (Co).  The checker of the opt-in fused kernel fsb_volume_stem."""
    vol = torch.cat([gwc_volume(left, right, maxdisp, num_groups), concat_volume(proj_left, proj_right, maxdisp)], 1)
    w = weight.reshape(weight.shape[0], -1)
    return torch.einsum("oc,bcdhw->bodhw", w, vol) + bias.reshape(1, -1, 1, 1, 1)


# --------------------------------------------------------------------------
# geometry-encoding volume  (core/geometry.py:17-77, core/utils/utils.py:44-55)
# --------------------------------------------------------------------------
def row_correlation(f1: Tensor, f2: Tensor) -> Tensor:
    """All-pairs per-row correlation (core/geometry.py:68-77).

    corr[b,h,w1,w2] = sum_c f1hat[b,c,h,w1] * f2hat[b,c,h,w2], features
    L2-normalised over all C channels, fp32.  Returned as (B,H,W1,W2); the
    reference adds a singleton axis before W2.
    """
    a = l2_normalize(f1.float(), 1)
    b = l2_normalize(f2.float(), 1)
    return torch.einsum("bchw,bchv->bhwv", a, b)


def halve_last(x: Tensor) -> Tensor:
    """avg_pool2d(kernel [1,2], stride [1,2]) over the last axis
    (core/geometry.py:35,39): mean of adjacent pairs, floor(len/2) outputs."""
    n = x.shape[-1] // 2
    return (x[..., 0 : 2 * n : 2] + x[..., 1 : 2 * n : 2]) * 0.5


def corr_pyramid(f1: Tensor, f2: Tensor, num_levels: int) -> List[Tensor]:
    """Levels of the init-correlation pyramid as (B*H*W1, W2 // 2**l)
    (core/geometry.py:25,31,38-40)."""
    c = row_correlation(f1, f2)
    B, H, W1, W2 = c.shape
    levels = [c.reshape(B * H * W1, W2)]
    for _ in range(num_levels - 1):
        levels.append(halve_last(levels[-1]))
    return levels


def geo_pyramid(volume: Tensor, num_levels: int) -> List[Tensor]:
    """Levels of the geometry-volume pyramid in the REFERENCE layout
    (B*H*W, Cg, D // 2**l) (core/geometry.py:28-36)."""
    B, Cg, D, H, W = volume.shape
    lv0 = volume.float().permute(0, 3, 4, 1, 2).reshape(B * H * W, Cg, D).contiguous()
    levels = [lv0]
    for _ in range(num_levels - 1):
        levels.append(halve_last(levels[-1]))
    return levels


def lerp_rows(rows: Tensor, x: Tensor) -> Tensor:
    """1-D linear interpolation of ``rows`` (N,C,Wl) at pixel positions
    ``x`` (N,K) -> (N,C,K).

    Restates bilinear_sampler (core/utils/utils.py:44-55) for the H == 1 case:
    the coordinate goes through 2x/(Wl-1)-1 and back (fp32 round trip, as
    grid_sample with align_corners=True does), each of the two taps is zeroed
    on its own when its index falls outside [0, Wl-1] (padding_mode='zeros').
    """
    N, C, Wl = rows.shape
    assert Wl > 1, "a pyramid level of length 1 divides by zero in the reference"
    xn = 2 * x / (Wl - 1) - 1
    ix = ((xn + 1) / 2) * (Wl - 1)
    i0 = torch.floor(ix)
    w_hi = ix - i0
    w_lo = (i0 + 1) - ix
    i0 = i0.long()
    i1 = i0 + 1

    def tap(idx: Tensor) -> Tensor:
        ok = (idx >= 0) & (idx <= Wl - 1)
        g = torch.gather(rows, 2, idx.clamp(0, Wl - 1)[:, None, :].expand(N, C, -1))
        return g * ok[:, None, :].to(rows.dtype)

    return tap(i0) * w_lo[:, None, :] + tap(i1) * w_hi[:, None, :]


def geo_lookup(
    geo_levels: Sequence[Tensor],
    corr_levels: Sequence[Tensor],
    disp: Tensor,
    coords: Tensor,
    radius: int,
) -> Tensor:
    """Per-iteration lookup (core/geometry.py:43-65).

    ``geo_levels[l]`` is (BHW, Cg, D_l) (reference layout), ``corr_levels[l]``
    is (BHW, W2_l).  Output (B, L*K*(Cg+1), H, W), channel order per level:
    geo channel c at offset k -> c*K + k, then the K correlation samples.
    """
    B, _, H, W = disp.shape
    N = B * H * W
    K = 2 * radius + 1
    dx = torch.linspace(-radius, radius, K, dtype=torch.float32).reshape(1, K)
    d = disp.float().reshape(N, 1)
    xc = coords.float().reshape(N, 1)
    out = []
    for l, (g, c) in enumerate(zip(geo_levels, corr_levels)):
        s = 2 ** l
        x_geo = dx + d / s
        out.append(lerp_rows(g, x_geo).reshape(N, -1))
        x_cor = xc / s - d / s + dx
        out.append(lerp_rows(c[:, None, :], x_cor).reshape(N, -1))
    out = torch.cat(out, dim=1)
    return out.reshape(B, H, W, -1).permute(0, 3, 1, 2).contiguous()


def corr_band_geometry(d_band: int, radius: int, num_levels: int) -> Tuple[int, int, int]:
    """(bw0, off, align) of the OPT-IN banded init correlation (include/fsb200.h, "BANDED form"): the pixel in image
    column w1 keeps the level-0 columns [s0, s0 + bw0), s0 = floor((w1 - off) / align) * align, i.e. what a lookup
    with disparities in [0, d_band] can touch at any level (core/geometry.py:57: x = (w1 - disp) / 2^l + dx)."""
    align = 4 << (num_levels - 1)
    left = (radius + 2) << (num_levels - 1)
    right = (radius + 3) << (num_levels - 1)
    bw0 = -(-(d_band + left + right + align - 1) // align) * align
    return bw0, d_band + left, align


def band_limit_corr_pyramid(corr_levels: Sequence[Tensor], W1: int, d_band: int, radius: int) -> List[Tensor]:
    """The band form's semantics restated on the full pyramid: every (BHW, W2_l) level with the entries OUTSIDE the
    pixel's band set to zero.  Not a reference behaviour -- the reference keeps the whole row -- so a lookup through
    this pyramid equals the reference's exactly when it stays inside the band."""
    L = len(corr_levels)
    bw0, off, align = corr_band_geometry(d_band, radius, L)
    out = []
    for l, c in enumerate(corr_levels):
        N, Wl = c.shape
        w1 = torch.arange(N) % W1
        s0 = torch.div(w1 - off, align, rounding_mode="floor") * align
        sl = torch.div(s0, 2 ** l, rounding_mode="floor")[:, None]
        col = torch.arange(Wl)[None, :]
        keep = (col >= sl) & (col < sl + (bw0 >> l))
        out.append(c * keep.to(c.dtype))
    return out


class GeoEncodingVolume:
    """Oracle counterpart of Combined_Geo_Encoding_Volume (core/geometry.py:17-65).  ``corr_band`` restates the
    opt-in banded form (entries outside the band zeroed, see band_limit_corr_pyramid)."""

    def __init__(self, f1: Tensor, f2: Tensor, volume: Tensor, num_levels: int = 2, radius: int = 4,
                 corr_band: int = 0):
        self.num_levels = num_levels
        self.radius = radius
        self.corr_levels = corr_pyramid(f1, f2, num_levels)
        if corr_band:
            self.corr_levels = band_limit_corr_pyramid(self.corr_levels, f1.shape[3], corr_band, radius)
        self.geo_levels = geo_pyramid(volume, num_levels)

    def __call__(self, disp: Tensor, coords: Tensor) -> Tensor:
        return geo_lookup(self.geo_levels, self.corr_levels, disp, coords, self.radius)


# --------------------------------------------------------------------------
# update block  (core/update.py:20-159, core/submodule.py:565-591)
# --------------------------------------------------------------------------
def _conv(p: Params, name: str, x: Tensor, pad: int, groups: int = 1) -> Tensor:
    return F.conv2d(x, p[name + ".weight"], p[name + ".bias"], padding=pad, groups=groups)


def pool2x(x: Tensor) -> Tensor:
    """core/update.py:72-73 (3x3 average, stride 2, pad 1, padded zeros counted)."""
    return F.avg_pool2d(x, 3, stride=2, padding=1)


def resize_like(x: Tensor, ref: Tensor) -> Tensor:
    """core/update.py:78-80 (bilinear, align_corners=True, to ref's H,W)."""
    return F.interpolate(x, ref.shape[2:], mode="bilinear", align_corners=True)


def raft_gru(p: Params, pre: str, h: Tensor, x: Tensor, hx: Tensor) -> Tensor:
    """RaftConvGRU.forward (core/update.py:90-95).  Kernel size is read from
    the weights (1x1 small, 3x3 large)."""
    pad = p[pre + ".convz.weight"].shape[-1] // 2
    z = torch.sigmoid(_conv(p, pre + ".convz", hx, pad))
    r = torch.sigmoid(_conv(p, pre + ".convr", hx, pad))
    q = torch.tanh(_conv(p, pre + ".convq", torch.cat([r * h, x], 1), pad))
    return (1 - z) * h + z * q


def selective_gru(p: Params, pre: str, att: Tensor, h: Tensor, *xs: Tensor) -> Tensor:
    """SelectiveConvGRU.forward (core/update.py:112-119)."""
    x = torch.cat(xs, 1)
    x = F.relu(_conv(p, pre + ".conv0.0", x, 1))
    hx = F.relu(_conv(p, pre + ".conv1.0", torch.cat([x, h], 1), 1))
    small = raft_gru(p, pre + ".small_gru", h, x, hx)
    large = raft_gru(p, pre + ".large_gru", h, x, hx)
    return small * att + large * (1 - att)


def motion_encoder(p: Params, disp: Tensor, corr: Tensor) -> Tensor:
    """BasicMotionEncoder.forward (core/update.py:62-70)."""
    c = F.relu(_conv(p, "encoder.convc1", corr, 0))
    c = F.relu(_conv(p, "encoder.convc2", c, 1))
    d = F.relu(_conv(p, "encoder.convd1", disp, 3))
    d = F.relu(_conv(p, "encoder.convd2", d, 1))
    o = F.relu(_conv(p, "encoder.conv", torch.cat([c, d], 1), 1))
    return torch.cat([o, disp], 1)


def _edgenext(p: Params, pre: str, x: Tensor) -> Tensor:
    """EdgeNextConvEncoder.forward with norm=None (core/submodule.py:579-591)."""
    C = x.shape[1]
    y = _conv(p, pre + ".dwconv", x, p[pre + ".dwconv.weight"].shape[-1] // 2, groups=C)
    y = y.permute(0, 2, 3, 1)
    y = F.linear(y, p[pre + ".pwconv1.weight"], p[pre + ".pwconv1.bias"])
    y = F.gelu(y)
    y = F.linear(y, p[pre + ".pwconv2.weight"], p[pre + ".pwconv2.bias"])
    y = p[pre + ".gamma"] * y
    return x + y.permute(0, 3, 1, 2)


def disp_head(p: Params, x: Tensor) -> Tensor:
    """DispHead.forward (core/update.py:20-32)."""
    y = F.relu(_conv(p, "disp_head.conv.0", x, 1))
    y = _edgenext(p, "disp_head.conv.2", y)
    y = _edgenext(p, "disp_head.conv.3", y)
    return _conv(p, "disp_head.conv.4", y, 1)


def mask_head(p: Params, x: Tensor) -> Tensor:
    """mask branch incl. the 0.25 scale (core/update.py:134-139,158)."""
    y = F.relu(_conv(p, "mask.0", x, 1))
    y = F.relu(_conv(p, "mask.2", y, 1))
    return 0.25 * y


def update_block(
    p: Params,
    net: List[Tensor],
    inp: Sequence[Tensor],
    corr: Tensor,
    disp: Tensor,
    att: Sequence[Tensor],
) -> Tuple[List[Tensor], Tensor, Tensor]:
    """BasicSelectiveMultiUpdateBlock.forward (core/update.py:141-159).  Mutates and returns ``net`` like the
    reference.  n_gru_layers is read off the parameter dict (which GRUs it holds, core/update.py:128-132):
    3 = gru16 + gru08 + gru04; 2 = gru08 (two inputs) + gru04; 1 = NO GRU at all -- the reference only calls
    gru04 under ``n_gru_layers > 1`` (:152), so net passes through and only the two heads run."""
    n = 3 if "gru16.conv0.0.weight" in p else 2 if "gru08.conv0.0.weight" in p else 1
    if n == 3:
        net[2] = selective_gru(p, "gru16", att[2], net[2], inp[2], pool2x(net[1]))
    if n >= 2:
        if n > 2:
            net[1] = selective_gru(p, "gru08", att[1], net[1], inp[1], pool2x(net[0]), resize_like(net[2], net[1]))
        else:
            net[1] = selective_gru(p, "gru08", att[1], net[1], inp[1], pool2x(net[0]))
        motion = torch.cat([inp[0], motion_encoder(p, disp, corr)], 1)
        net[0] = selective_gru(p, "gru04", att[0], net[0], motion, resize_like(net[1], net[0]))
    delta = disp_head(p, net[0])
    mask = mask_head(p, net[0])
    return net, mask, delta


# --------------------------------------------------------------------------
# the whole refinement loop  (core/foundation_stereo.py:229-248)
# --------------------------------------------------------------------------
def refine(
    p: Params,
    geo: GeoEncodingVolume,
    net: List[Tensor],
    inp: Sequence[Tensor],
    att: Sequence[Tensor],
    disp: Tensor,
    iters: int,
) -> Tuple[Tensor, Tensor, List[Tensor]]:
    """disp_{t+1} = disp_t + delta(lookup(disp_t)); returns the final 1/4-res
    disparity, the last mask features and the final hidden states."""
    B, _, H, W = disp.shape
    coords = torch.arange(W, dtype=torch.float32).reshape(1, 1, W, 1).repeat(B, H, 1, 1)
    disp = disp.float()
    net = list(net)
    mask = None
    for _ in range(iters):
        feat = geo(disp, coords)
        net, mask, delta = update_block(p, net, inp, feat, disp, att)
        disp = disp + delta.float()
    return disp, mask, net


# --------------------------------------------------------------------------
# the steps either side of the loop  (core/submodule.py:431-435, 456-468;
# core/foundation_stereo.py:183-191, 218-220)
# --------------------------------------------------------------------------
def disparity_regression(prob: Tensor) -> Tensor:
    """Expected disparity (core/submodule.py:431-435): (B,D,H,W) -> (B,1,H,W),
    out = sum_d prob[:,d] * d in the dtype of ``prob``."""
    D = prob.shape[1]
    d = torch.arange(D, dtype=prob.dtype).reshape(1, D, 1, 1)
    return (prob * d).sum(1, keepdim=True)


def softmax_disparity_regression(logits: Tensor) -> Tensor:
    """core/foundation_stereo.py:218-220: softmax over the disparity axis, then
    the expected disparity; fp32 (softmax runs in fp32 under autocast)."""
    return disparity_regression(torch.softmax(logits.float(), 1))


def context_upsample(disp_low: Tensor, weights: Tensor) -> Tensor:
    """Convex 4x upsampling (core/submodule.py:456-468): every full-resolution
    pixel (Y,X) is a weighted sum of the 3x3 neighbourhood (zero padded, row-major
    k = ky*3+kx) of its low-resolution pixel (Y//4, X//4).  (b,1,h,w),(b,9,4h,4w)
    -> (b,4h,4w)."""
    b, _, h, w = disp_low.shape
    pad = F.pad(disp_low[:, 0], (1, 1, 1, 1))
    out = torch.zeros((b, 4 * h, 4 * w), dtype=torch.result_type(disp_low, weights))
    for ky in range(3):
        for kx in range(3):
            nb = pad[:, ky : ky + h, kx : kx + w]
            nb = nb.repeat_interleave(4, 1).repeat_interleave(4, 2)
            out = out + nb * weights[:, ky * 3 + kx]
    return out


def upsample_disp_tail(disp: Tensor, spx_logits: Tensor) -> Tensor:
    """Tail of FoundationStereo.upsample_disp (core/foundation_stereo.py:188-191):
    softmax over the 9 weights, disparity scaled by 4, convex upsampling -> (b,1,4h,4w) fp32."""
    return context_upsample(disp.float() * 4.0, torch.softmax(spx_logits.float(), 1)).unsqueeze(1).float()
