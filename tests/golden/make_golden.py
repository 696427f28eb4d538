#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by RUNNING THE REFERENCE.

Runs only in the build container, where /root/reference exists; the GPU box
never executes this file, it only reads the committed .npz fixtures.

    python tests/golden/make_golden.py

The reference's hot-path modules are imported unmodified after stubbing the
six third-party packages its Utils.py imports at module top but this image
lacks.  Inputs come from ``oracle.inputs`` / seeded
generators, weights from ``oracle.params`` loaded into the reference module
with strict=True.
"""
import os
import sys
import types
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("FS_REFERENCE", "/root/reference")


def import_reference():
    for name in ("trimesh", "open3d", "imageio", "transformations", "timm", "omegaconf"):
        sys.modules.setdefault(name, types.ModuleType(name)).__dict__.setdefault("__all__", [])
    if REF not in sys.path:
        sys.path.insert(0, REF)
    warnings.filterwarnings("ignore")
    from core.submodule import build_gwc_volume, build_concat_volume
    from core.geometry import Combined_Geo_Encoding_Volume
    from core.update import BasicSelectiveMultiUpdateBlock
    return build_gwc_volume, build_concat_volume, Combined_Geo_Encoding_Volume, BasicSelectiveMultiUpdateBlock


def save(name, **arrays):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **{k: (v.numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in arrays.items()})
    print(f"{name}.npz  {os.path.getsize(path) / 1024:.1f} KiB")


def ragged_disp(g, B, H, W, D):
    d = torch.rand(B, 1, H, W, generator=g) * D
    u = torch.rand(B, 1, H, W, generator=g)
    d = torch.where(u < 0.1, -6.0 * torch.rand(B, 1, H, W, generator=g), d)
    d = torch.where(u > 0.9, D + 6.0 * torch.rand(B, 1, H, W, generator=g), d)
    # a few exact integers and exact boundaries
    d.view(-1)[0] = 0.0
    d.view(-1)[1] = float(D - 1)
    d.view(-1)[2] = 3.0
    d.view(-1)[3] = -1.0
    return d


def main():
    torch.manual_seed(0)
    torch.set_num_threads(4)
    gwc_ref, cat_ref, GeoRef, UpdRef = import_reference()

    # ---- volumes: odd W, D > some widths, a zero feature vector (eps branch)
    g = torch.Generator().manual_seed(11)
    B, C, G, H, W, D, Cp = 2, 24, 8, 3, 23, 13, 5
    fl = torch.relu(torch.randn(B, C, H, W, generator=g))
    fr = torch.relu(torch.randn(B, C, H, W, generator=g))
    fl[0, :, 1, 4] = 0.0
    fr[1, :, 2, 7] = 0.0
    fr[0, 3:6, 0, 0] = 0.0  # one whole group (cpg=3) zero on the right
    pl = torch.randn(B, Cp, H, W, generator=g)
    pr = torch.randn(B, Cp, H, W, generator=g)
    save("volumes_f32", fl=fl, fr=fr, pl=pl, pr=pr, G=G, D=D,
         gwc=gwc_ref(fl, fr, D, G), cat=cat_ref(pl, pr, D))
    # fp16 storage, fp32 math (autocast behaviour)
    save("volumes_f16", fl=fl.half(), fr=fr.half(), pl=pl.half(), pr=pr.half(), G=G, D=D,
         gwc=gwc_ref(fl.half(), fr.half(), D, G), cat=cat_ref(pl.half(), pr.half(), D))
    # D larger than W
    Wn, Dn = 6, 9
    save("volumes_narrow", fl=fl[..., :Wn], fr=fr[..., :Wn], pl=pl[..., :Wn], pr=pr[..., :Wn], G=G, D=Dn,
         gwc=gwc_ref(fl[..., :Wn].contiguous(), fr[..., :Wn].contiguous(), Dn, G),
         cat=cat_ref(pl[..., :Wn].contiguous(), pr[..., :Wn].contiguous(), Dn))

    # ---- geometry-encoding volume: pyramids and lookup
    for tag, (B, C, H, W, D, Cg, L, r) in {
        "geo_odd": (1, 24, 4, 23, 13, 6, 3, 4),
        "geo_std": (2, 32, 3, 32, 24, 28, 2, 4),
        "geo_l4": (1, 16, 2, 40, 32, 28, 4, 3),
    }.items():
        g = torch.Generator().manual_seed(len(tag) * 7 + 1)
        f1 = torch.relu(torch.randn(B, C, H, W, generator=g))
        f2 = torch.relu(torch.randn(B, C, H, W, generator=g))
        f1[0, :, 0, 1] = 0.0
        vol = torch.randn(B, Cg, D, H, W, generator=g)
        disp = ragged_disp(g, B, H, W, D)
        coords = torch.arange(W, dtype=torch.float32).reshape(1, 1, W, 1).repeat(B, H, 1, 1)
        dx = torch.linspace(-r, r, 2 * r + 1).reshape(1, 1, 2 * r + 1, 1)
        fn = GeoRef(f1, f2, vol, num_levels=L, dx=dx)
        arrays = dict(f1=f1, f2=f2, volume=vol, disp=disp, coords=coords, L=L, r=r, lookup=fn(disp, coords))
        for l in range(L):
            arrays[f"geo{l}"] = fn.geo_volume_pyramid[l].squeeze(2)
            arrays[f"corr{l}"] = fn.init_corr_pyramid[l].reshape(B * H * W, -1)
        save(tag, **arrays)

    # ---- update block: one teacher-forced call, then a 3-iteration free run
    from oracle.params import make_update_block_params
    from oracle.inputs import make_inputs
    L, r, Cg = 2, 4, 28
    args = types.SimpleNamespace(corr_levels=L, corr_radius=r, n_gru_layers=3)
    ub = UpdRef(args, 128, volume_dim=Cg).eval()
    params = make_update_block_params(seed=3, corr_levels=L, corr_radius=r, volume_dim=Cg)
    ub.load_state_dict(params, strict=True)
    x = make_inputs(B=1, C=32, H=8, W=12, D=10, Cg=Cg, seed=5)
    dx = torch.linspace(-r, r, 2 * r + 1).reshape(1, 1, 2 * r + 1, 1)
    with torch.no_grad():
        fn = GeoRef(x["fl"], x["fr"], x["volume"], num_levels=L, dx=dx)
        corr = fn(x["disp"], x["coords"])
        net = [t.clone() for t in x["net"]]
        net, mask, delta = ub(net, x["inp"], corr, x["disp"], x["att"])
        save("update_step", seed_params=3, seed_inputs=5, corr=corr, net0=net[0], net1=net[1], net2=net[2],
             mask=mask, delta=delta)
        # free-running loop, as core/foundation_stereo.py:236-242
        net = [t.clone() for t in x["net"]]
        disp = x["disp"].clone()
        trace = []
        for _ in range(3):
            feat = fn(disp, x["coords"])
            net, mask, delta = ub(net, x["inp"], feat, disp, x["att"])
            disp = disp + delta.float()
            trace.append(disp.clone())
        save("refine_3it", seed_params=3, seed_inputs=5, disp1=trace[0], disp2=trace[1], disp3=trace[2],
             mask=mask, net0=net[0])

    # ---- the steps either side of the loop: disparity_regression, context_upsample (+ their softmaxes)
    import torch.nn.functional as F
    from core.submodule import disparity_regression, context_upsample
    g = torch.Generator().manual_seed(23)
    B, D, H, W = 2, 13, 5, 7          # odd H*W: the scalar path
    logits = 3.0 * torch.randn(B, D, H, W, generator=g)
    logits[0, :, 0, 0] = 0.0          # uniform distribution
    logits[0, 5, 0, 1] = 60.0         # one-hot
    prob = F.softmax(logits, 1)
    B2, D2, H2, W2 = 1, 24, 4, 8      # H*W % 4 == 0: the vector path
    logits2 = 2.0 * torch.randn(B2, D2, H2, W2, generator=g)
    prob2 = F.softmax(logits2, 1)
    h, w = 3, 5
    disp = 20.0 * torch.rand(B, 1, h, w, generator=g)
    spx = 2.0 * torch.randn(B, 9, 4 * h, 4 * w, generator=g)
    wts = F.softmax(spx, 1)
    save("post", logits=logits, prob=prob, reg=disparity_regression(prob, D),
         logits2=logits2, prob2=prob2, reg2=disparity_regression(prob2, D2),
         reg2_f16=disparity_regression(prob2.half(), D2),
         disp=disp, spx=spx, wts=wts, up=context_upsample(disp, wts),
         up_tail=context_upsample(disp * 4., wts).unsqueeze(1))


if __name__ == "__main__":
    main()
