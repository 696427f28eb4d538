"""Drop-in replacement for core/geometry.py's Combined_Geo_Encoding_Volume.

Constructor and call signatures, attribute names and results follow
core/geometry.py:17-77.  Construction runs kernel 2 (init correlation + pooled
levels) and kernel 2b (volume transpose + pooled levels); each call runs kernel 3
(all levels, both volumes, NCHW output) -- one launch and no host sync instead of
the reference's ~70 launches and 2*L ``torch.unique`` syncs per iteration
(core/utils/utils.py:49).
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import _lib, ops


class Combined_Geo_Encoding_Volume:
    def __init__(self, init_fmap1: torch.Tensor, init_fmap2: torch.Tensor, geo_volume: torch.Tensor,
                 num_levels: int = 2, dx: Optional[torch.Tensor] = None, precision: int = _lib.FSB_PREC_TF32X3,
                 workspace: Optional[dict] = None, nhwc16_out=None, corr_band: Optional[int] = None):
        """``workspace``: optional {"corr": [...], "geo": [...]} caller-owned level buffers of the
        right shapes (stable addresses, so a CUDA graph that holds the lookup can be replayed
        for every pair).  ``nhwc16_out``: optional callable (B,H,W) -> 16-bit (B,H,W,Cpad)
        buffer; when given, a call writes the fused 16-bit NHWC layout into that buffer and
        returns its (B, L*K*(Cg+1), H, W) channels-last view (what SelectiveUpdateBlock
        consumes in place) instead of the reference's fp32 NCHW tensor.
        ``corr_band``: opt-in BANDED init correlation for disparities in [0, corr_band] (e.g. max_disp // 4): only the
        band of the W x W matrix a lookup in that range can touch is computed and stored ((BHW, bw0 >> l) per level,
        include/fsb200.h).  Inside the band results are identical to the full form; a disparity outside it reads zeros
        where the reference reads real correlations (core/geometry.py:57 never clamps) -- hence opt-in."""
        if dx is None:
            raise ValueError("dx (the (1,1,2r+1,1) offsets, core/foundation_stereo.py:177-179) is required")
        K = dx.numel()
        r = (K - 1) // 2
        if K != 2 * r + 1 or r < 1:
            raise ValueError(f"dx must hold the 2r+1 integer offsets -r..r, got {K} values")
        if not dx.is_cuda:  # CPU tensor in the reference model: checking costs no device sync
            want = torch.linspace(-r, r, K)
            if not torch.equal(dx.reshape(-1).float().cpu(), want):
                raise ValueError("dx must be linspace(-r, r, 2r+1): the kernel samples integer offsets")
        self.num_levels = int(num_levels)
        self.radius = r
        self.dx = dx
        self.precision = precision
        B, Cg, D, H, W = geo_volume.shape
        assert init_fmap1.shape[0] == B and init_fmap1.shape[2] == H and init_fmap1.shape[3] == W
        self._shape = (B, Cg, D, H, W)
        self._w2 = init_fmap2.shape[3]
        # lookup layout: geo level l is (BHW, D>>l, Cg), corr level l is (BHW, W2>>l)
        self._nhwc16_out = nhwc16_out
        self.corr_band = int(corr_band) if corr_band else None
        if self.corr_band is not None and precision == _lib.FSB_PREC_FP32:
            raise ValueError("the banded correlation exists on the tensor-core path (FSB_PREC_TF32X3 / FSB_PREC_TF32)")
        if workspace is not None:
            self.corr_levels = list(workspace["corr"])
            self.geo_levels = list(workspace["geo"])
            if self.corr_band is not None:
                ops.corr_pyramid_band(init_fmap1, init_fmap2, self.num_levels, precision, self.corr_band, r, self.corr_levels)
            else:
                ops.corr_pyramid_into(init_fmap1, init_fmap2, self.corr_levels, precision)
            ops.geo_pyramid_into(geo_volume, self.geo_levels)
        else:
            if self.corr_band is not None:
                self.corr_levels = list(ops.corr_pyramid_band(init_fmap1, init_fmap2, self.num_levels, precision,
                                                              self.corr_band, r))
            else:
                self.corr_levels = list(ops.corr_pyramid(init_fmap1, init_fmap2, self.num_levels, precision))
            self.geo_levels: List[torch.Tensor] = list(ops.geo_pyramid(geo_volume, self.num_levels))

    @staticmethod
    def make_workspace(B: int, Cg: int, D: int, H: int, W: int, W2: int, num_levels: int, device,
                       corr_band: Optional[int] = None, radius: int = 4) -> dict:
        """Level buffers in the lookup layout: geo (BHW, D>>l, Cg), corr (BHW, W2>>l) -- or (BHW, bw0>>l) for a
        banded correlation -- fp32."""
        n = B * H * W
        if corr_band:
            W2 = ops.corr_band_geometry(corr_band, radius, num_levels)[0]
        return {"corr": [torch.empty((n, W2 >> l), dtype=torch.float32, device=device) for l in range(num_levels)],
                "geo": [torch.empty((n, D >> l, Cg), dtype=torch.float32, device=device) for l in range(num_levels)]}

    # -- the reference's attribute names, as views in the reference's shapes -------------
    @property
    def geo_volume_pyramid(self) -> List[torch.Tensor]:
        """[(BHW, Cg, 1, D_l)] like core/geometry.py:32-36 (non-contiguous views)."""
        return [g.permute(0, 2, 1).unsqueeze(2) for g in self.geo_levels]

    @property
    def init_corr_pyramid(self) -> List[torch.Tensor]:
        """[(BHW, 1, 1, W2_l)] like core/geometry.py:31-40 (the stored band, (BHW, 1, 1, bw0 >> l), in band mode)."""
        return [c[:, None, None, :] for c in self.corr_levels]

    def __call__(self, disp: torch.Tensor, coords: torch.Tensor, low_memory: bool = False) -> torch.Tensor:
        """(B,1,H,W) disparity, (B,H,W,1) x coordinates -> (B, L*K*(Cg+1), H, W) fp32.
        ``low_memory`` is accepted and ignored like the reference (core/utils/utils.py:44)."""
        if self._nhwc16_out is not None:
            return self.lookup_nhwc16(disp, coords, self._nhwc16_out(*[disp.shape[i] for i in (0, 2, 3)]))
        if self.corr_band is not None:
            return ops.geo_lookup_band(self.geo_levels, self.corr_levels, disp, coords, self.radius, self.corr_band)
        return ops.geo_lookup(self.geo_levels, self.corr_levels, disp, coords, self.radius)

    def lookup_nhwc16(self, disp: torch.Tensor, coords: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """Fused-output lookup into the 16-bit (B,H,W,Cpad) buffer ``out``; returns the
        (B, L*K*(Cg+1), H, W) view of it (channels-last strides, pad channels hidden)."""
        B, H, W, Cpad = out.shape
        if self.corr_band is not None:
            ops.geo_lookup_band(self.geo_levels, self.corr_levels, disp, coords, self.radius, self.corr_band, out)
        else:
            ops.geo_lookup_nhwc16(self.geo_levels, self.corr_levels, disp.float().contiguous(),
                                  coords.float().contiguous(), self.radius, out)
        planes = self.num_levels * (2 * self.radius + 1) * (self._shape[1] + 1)
        return torch.as_strided(out, (B, planes, H, W), (H * W * Cpad, 1, W * Cpad, Cpad))

    @staticmethod
    def corr(fmap1: torch.Tensor, fmap2: torch.Tensor) -> torch.Tensor:
        """(B,C,H,W1),(B,C,H,W2) -> (B,H,W1,1,W2) fp32 (core/geometry.py:68-77)."""
        B, _, H, W1 = fmap1.shape
        W2 = fmap2.shape[3]
        (c,) = ops.corr_pyramid(fmap1, fmap2, 1, _lib.FSB_PREC_TF32X3)
        return c.reshape(B, H, W1, 1, W2)
