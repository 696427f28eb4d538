"""Pins the CPU oracle to the reference's own outputs (tests/golden/*.npz,
made by tests/golden/make_golden.py which imports and runs /root/reference)."""
import pytest
import torch

import oracle
from conftest import load_golden, rel_err


@pytest.mark.parametrize("name", ["volumes_f32", "volumes_f16", "volumes_narrow"])
def test_volumes(name):
    g = load_golden(name)
    gwc = oracle.gwc_volume(g["fl"], g["fr"], g["D"], g["G"])
    cat = oracle.concat_volume(g["pl"], g["pr"], g["D"])
    assert gwc.dtype == g["gwc"].dtype and gwc.shape == g["gwc"].shape
    assert torch.equal(cat, g["cat"])  # pure copy: bit exact
    tol = 1e-6 if gwc.dtype == torch.float32 else 1e-3
    assert rel_err(gwc, g["gwc"]) <= tol
    # the zero region w < d must be exactly zero
    D, W = gwc.shape[2], gwc.shape[4]
    for d in range(D):
        assert gwc[:, :, d, :, : min(d, W)].abs().max().item() == 0 if d > 0 else True


@pytest.mark.parametrize("name", ["geo_odd", "geo_std", "geo_l4"])
def test_geo_encoding(name):
    g = load_golden(name)
    L, r = g["L"], g["r"]
    vol = oracle.GeoEncodingVolume(g["f1"], g["f2"], g["volume"], L, r)
    for l in range(L):
        assert vol.geo_levels[l].shape == g[f"geo{l}"].shape
        assert torch.equal(vol.geo_levels[l], g[f"geo{l}"])  # copy + exact halving
        assert rel_err(vol.corr_levels[l], g[f"corr{l}"]) <= 2e-6
    out = vol(g["disp"], g["coords"])
    assert out.shape == g["lookup"].shape
    assert rel_err(out, g["lookup"]) <= 1e-5


def test_update_block_step():
    g = load_golden("update_step")
    p = oracle.make_update_block_params(seed=g["seed_params"], corr_levels=2, corr_radius=4, volume_dim=28)
    x = oracle.make_inputs(B=1, C=32, H=8, W=12, D=10, Cg=28, seed=g["seed_inputs"])
    net = [t.clone() for t in x["net"]]
    with torch.no_grad():
        net, mask, delta = oracle.update_block(p, net, x["inp"], g["corr"], x["disp"], x["att"])
    for i in range(3):
        assert rel_err(net[i], g[f"net{i}"]) <= 1e-5
    assert rel_err(mask, g["mask"]) <= 1e-5
    assert rel_err(delta, g["delta"]) <= 1e-5


def test_refine_loop():
    g = load_golden("refine_3it")
    p = oracle.make_update_block_params(seed=g["seed_params"], corr_levels=2, corr_radius=4, volume_dim=28)
    x = oracle.make_inputs(B=1, C=32, H=8, W=12, D=10, Cg=28, seed=g["seed_inputs"])
    with torch.no_grad():
        geo = oracle.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], 2, 4)
        d1, _, _ = oracle.refine(p, geo, x["net"], x["inp"], x["att"], x["disp"], 1)
        d3, mask, net = oracle.refine(p, geo, x["net"], x["inp"], x["att"], x["disp"], 3)
    assert (d1 - g["disp1"]).abs().max().item() <= 1e-4
    assert (d3 - g["disp3"]).abs().max().item() <= 1e-3
    assert rel_err(net[0], g["net0"]) <= 1e-3


def test_param_shapes_match_reference_listing():
    # the reference module has 16 348 640 parameters (L=2, r=4, Cg=28)
    shapes = oracle.update_block_param_shapes(2, 4, 128, 28)
    n = sum(int(torch.tensor(s).prod()) for s in shapes.values())
    assert n == 16_348_640


def test_post_steps():
    """disparity_regression / context_upsample and their fused-softmax forms against the reference's
    outputs (core/submodule.py:431-435, 456-468; core/foundation_stereo.py:188-189, 218-220)."""
    g = load_golden("post")
    assert torch.equal(oracle.disparity_regression(g["prob"]), g["reg"])
    assert torch.equal(oracle.disparity_regression(g["prob2"]), g["reg2"])
    r16 = oracle.disparity_regression(g["prob2"].half())
    assert r16.dtype == torch.float16 and rel_err(r16, g["reg2_f16"]) <= 2e-3
    assert rel_err(oracle.softmax_disparity_regression(g["logits"]), g["reg"]) <= 1e-6
    up = oracle.context_upsample(g["disp"], g["wts"])
    assert up.shape == g["up"].shape and rel_err(up, g["up"]) <= 1e-6
    tail = oracle.upsample_disp_tail(g["disp"], g["spx"])
    assert tail.shape == g["up_tail"].shape and rel_err(tail, g["up_tail"]) <= 1e-6
