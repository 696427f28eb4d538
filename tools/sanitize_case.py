#!/usr/bin/env python
"""One small pass over every kernel of libfsb200.so for compute-sanitizer (tools/sanitize.sh): volumes, init correlation
(TMA-fed full + band forms, register-transpose form, CUDA-core form), geometry pyramid, both lookups, one update-block
step (pair mode, row-reuse mode, gate epilogues, fused MLP, depthwise / 7x7 / dot-product helpers) and the two post
kernels.  No graphs (the sanitizer instruments plain launches), tiny shapes, results compared with the CPU oracle so a
corrupted run cannot pass silently."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import foundationstereo_b200 as fsb
import oracle
from foundationstereo_b200 import ops
from foundationstereo_b200._lib import FSB_PREC_FP32, FSB_PREC_TF32, FSB_PREC_TF32X3

torch.manual_seed(0)
B, C, H, W, D, L, r = 2, 32, 16, 64, 16, 2, 4
x = oracle.make_inputs(B=B, C=C, H=H, W=W, D=D, Cg=28, seed=3)
c = lambda t: t.cuda()
comb = fsb.build_combined_volume(c(x["fl"]), c(x["fr"]), c(x["pl"]), c(x["pr"]), D, 8)
gwc = fsb.build_gwc_volume(c(x["fl"]).half(), c(x["fr"]).half(), D, 8)
dx = torch.linspace(-r, r, 2 * r + 1).reshape(1, 1, -1, 1)
for prec in (FSB_PREC_TF32X3, FSB_PREC_TF32, FSB_PREC_FP32):
    ops.corr_pyramid(c(x["fl"]), c(x["fr"]), L, prec)
ops.corr_pyramid(c(x["fl"])[..., :62].contiguous(), c(x["fr"])[..., :62].contiguous(), L, FSB_PREC_TF32X3)  # W % 4 != 0: corr_umma
full = fsb.Combined_Geo_Encoding_Volume(c(x["fl"]), c(x["fr"]), c(x["volume"]), num_levels=L, dx=dx)
band = fsb.Combined_Geo_Encoding_Volume(c(x["fl"]), c(x["fr"]), c(x["volume"].half()), num_levels=L, dx=dx, corr_band=D)
feat = full(c(x["disp"]), c(x["coords"]))
featb = band(c(x["disp"]).clamp(0, D), c(x["coords"]))
p = oracle.make_update_block_params(seed=1)
ub = fsb.SelectiveUpdateBlock(p, "cuda", use_graph=False)
net, mask, delta = ub([c(t) for t in x["net"]], [c(t) for t in x["inp"]], feat, c(x["disp"]), [c(t) for t in x["att"]])
net, mask, delta = [t.clone() for t in net], mask.clone(), delta.clone()   # views of the plan's buffers: refine() below reuses them
ws = fsb.Combined_Geo_Encoding_Volume.make_workspace(B, 28, D, H, W, W, L, "cuda")
geo = fsb.Combined_Geo_Encoding_Volume(c(x["fl"]), c(x["fr"]), c(x["volume"]), num_levels=L, dx=dx, workspace=ws)
disp2, mask2, _ = ub.refine(geo, [c(t) for t in x["net"]], [c(t) for t in x["inp"]], [c(t) for t in x["att"]], c(x["disp"]), 2)
reg = fsb.softmax_disparity_regression(torch.randn(B, D, H, W, device="cuda"))
up = fsb.softmax_context_upsample(disp2, torch.randn(B, 9, 4 * H, 4 * W, device="cuda"), 4.0)
torch.cuda.synchronize()
with torch.no_grad():
    g = oracle.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], L, r)
    want_feat = g(x["disp"], x["coords"])
    wn, wm, wd = oracle.update_block(p, [t.clone() for t in x["net"]], x["inp"], want_feat, x["disp"], x["att"])
rel = lambda a, b: float((a.double().cpu() - b.double()).abs().max() / b.double().abs().max())
assert rel(comb[:, :8], oracle.gwc_volume(x["fl"], x["fr"], D, 8)) <= 1e-4
assert rel(feat, want_feat) <= 1e-4
assert rel(delta, wd) <= 2e-2 and rel(net[0], wn[0]) <= 2e-2 and rel(mask, wm) <= 2e-2
print("sanitize case ok:", fsb.launch_count(), "launches")
