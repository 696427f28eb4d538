"""BASELINE.json configs[2..4] as parity-test cases (configs[1] is covered in test_gpu_parity.py and is
the bench line): full-size, size-independent properties of kernels 1-3 checked against plain
torch fp32 ops on the GPU (TF32 disabled), plus one update-block step against the CPU oracle at
the KITTI shape.  Shapes from SURVEY.md section 8 (1/4 resolution of the padded image)."""
import pytest
import torch

import oracle
from conftest import rel_err

pytestmark = pytest.mark.gpu

CONFIGS = {
    # name: (B, C, H, W, D)
    "vitl_hier_pass_a_512x768": (1, 224, 128, 192, 104),
    "vitl_hier_pass_b_1024x1536": (1, 224, 256, 384, 104),
    "kitti_375x1242_batch2": (2, 128, 96, 312, 104),
    "erp_1024x2048": (1, 128, 256, 512, 104),
}


@pytest.fixture(scope="module")
def fsb():
    import foundationstereo_b200 as f
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    return f


@pytest.mark.parametrize("name", list(CONFIGS))
def test_config_shapes(fsb, name):
    B, C, H, W, D = CONFIGS[name]
    G, Cp, Cg, L, r = 8, 12, 28, 2, 4
    K = 2 * r + 1
    g = torch.Generator(device="cuda").manual_seed(3)
    rn = lambda *s: torch.randn(*s, generator=g, device="cuda")
    fl, fr = torch.relu(rn(B, C, H, W)), torch.relu(rn(B, C, H, W))
    pl, pr = rn(B, Cp, H, W), rn(B, Cp, H, W)
    # ---- kernel 1: one launch for both volumes
    comb = fsb.build_combined_volume(fl, fr, pl, pr, D, G)
    assert comb.shape == (B, G + 2 * Cp, D, H, W)
    gwc, cat = comb[:, :G], comb[:, G:]
    assert torch.equal(cat[:, :Cp], pl[:, :, None].expand(B, Cp, D, H, W))
    ln = torch.nn.functional.normalize(fl.reshape(B, G, C // G, H, W), dim=2)
    rnm = torch.nn.functional.normalize(fr.reshape(B, G, C // G, H, W), dim=2)
    for dd in (0, 1, D // 2, D - 1):
        assert torch.equal(cat[:, Cp:, dd, :, dd:], pr[..., : W - dd])
        assert cat[:, Cp:, dd, :, :dd].abs().sum().item() == 0.0
        assert gwc[:, :, dd, :, :dd].abs().sum().item() == 0.0
        want = (ln[..., dd:] * rnm[..., : W - dd]).sum(2)
        assert rel_err(gwc[:, :, dd, :, dd:], want) <= 1e-4
    del comb, gwc, cat, ln, rnm
    # ---- kernels 2, 2b
    volume = rn(B, Cg, D, H, W)
    dx = torch.linspace(-r, r, K).reshape(1, 1, -1, 1)
    vol = fsb.Combined_Geo_Encoding_Volume(fl, fr, volume, num_levels=L, dx=dx)
    c0, g0 = vol.corr_levels[0], vol.geo_levels[0]
    a = torch.nn.functional.normalize(fl, dim=1)
    b = torch.nn.functional.normalize(fr, dim=1)
    want = torch.einsum("bchw,bchv->bhwv", a, b).reshape(B * H * W, W)
    assert rel_err(c0, want) <= 1e-4
    del want, a, b
    assert torch.equal(vol.corr_levels[1], (c0[:, 0::2] + c0[:, 1::2]) * 0.5)
    assert torch.equal(g0.reshape(B, H, W, D, Cg).permute(0, 4, 3, 1, 2), volume)
    assert torch.equal(vol.geo_levels[1], (g0[:, 0::2] + g0[:, 1::2]) * 0.5)
    # ---- kernel 3: integer in-range disparities make the lookup a gather; fp32 NCHW and fused 16-bit
    # NHWC layouts hold the same samples
    disp = 2.0 * torch.randint(r, (D - r - 1) // 2, (B, 1, H, W), generator=g, device="cuda").float()
    coords = torch.arange(W, dtype=torch.float32, device="cuda").reshape(1, 1, W, 1).repeat(B, H, 1, 1)
    out = vol(disp, coords)
    assert out.shape == (B, L * K * (Cg + 1), H, W)
    idx = disp.reshape(-1, 1).long() + torch.arange(-r, r + 1, device="cuda")[None, :]
    gat = torch.gather(g0.permute(0, 2, 1), 2, idx[:, None, :].expand(-1, Cg, -1))
    got = out[:, : Cg * K].permute(0, 2, 3, 1).reshape(-1, Cg, K)
    assert rel_err(got, gat) <= 1e-4
    buf = torch.empty(B, H, W, 576, dtype=torch.float16, device="cuda")
    view = vol.lookup_nhwc16(disp, coords, buf)
    assert ((view.float() - out).abs() <= 2.0 ** -10 * out.abs() + 1e-6).all()


def test_update_block_step_kitti_shape(fsb):
    """One BasicSelectiveMultiUpdateBlock step at 96x312 (ragged pixel tiles at every pyramid level)
    against the fp32 CPU oracle; 16-bit operands / fp32 accumulation, tolerance 2e-2."""
    H, W = 96, 312
    p = oracle.make_update_block_params(seed=5)
    x = oracle.make_inputs(B=1, C=16, H=H, W=W, D=16, Cg=28, seed=12)
    with torch.no_grad():
        geo = oracle.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], 2, 4)
        corr = geo(x["disp"], x["coords"])
        want_net, want_mask, want_delta = oracle.update_block(p, [t.clone() for t in x["net"]], x["inp"], corr,
                                                              x["disp"], x["att"])
    ub = fsb.SelectiveUpdateBlock(p, "cuda")
    c = lambda t: t.cuda()
    net, mask, delta = ub([c(t) for t in x["net"]], [c(t) for t in x["inp"]], c(corr), c(x["disp"]),
                          [c(t) for t in x["att"]])
    for i in range(3):
        assert rel_err(net[i].cpu(), want_net[i]) <= 2e-2
    assert rel_err(mask.cpu(), want_mask) <= 2e-2
    assert rel_err(delta.cpu(), want_delta) <= 2e-2
