"""GPU parity of the steps either side of the refinement loop, through the C ABI:
disparity_regression (core/submodule.py:431-435) and context_upsample (core/submodule.py:456-468), plain
and with the softmaxes of core/foundation_stereo.py:218 / :188 fused in.  fp32 tolerance 1e-4 relative
(max|a-b| / max|b|), 16-bit inputs 2e-3 (fp16) / 1.6e-2 (bf16) as for the volumes."""
import pytest
import torch

import oracle
from conftest import load_golden, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope="module")
def fsb():
    import foundationstereo_b200 as f
    assert f.lib().fsb_version() >= 120
    return f


def test_post_golden(fsb):
    g = load_golden("post")
    n0 = fsb.launch_count()
    reg = fsb.disparity_regression(g["prob"].cuda(), g["prob"].shape[1])          # odd H*W: scalar kernel
    reg2 = fsb.disparity_regression(g["prob2"].cuda(), g["prob2"].shape[1])       # vector kernel
    assert fsb.launch_count() == n0 + 2
    assert reg.shape == g["reg"].shape and reg.dtype == torch.float32
    assert rel_err(reg.cpu(), g["reg"]) <= TOL and rel_err(reg2.cpu(), g["reg2"]) <= TOL
    r16 = fsb.disparity_regression(g["prob2"].half().cuda(), g["prob2"].shape[1])
    assert r16.dtype == torch.float16 and rel_err(r16.cpu(), g["reg2_f16"]) <= 2e-3
    assert rel_err(fsb.softmax_disparity_regression(g["logits"].cuda()).cpu(), g["reg"]) <= TOL
    assert rel_err(fsb.softmax_disparity_regression(g["logits2"].cuda()).cpu(), g["reg2"]) <= TOL
    up = fsb.context_upsample(g["disp"].cuda(), g["wts"].cuda())
    assert up.shape == g["up"].shape and up.dtype == torch.float32
    assert rel_err(up.cpu(), g["up"]) <= TOL
    tail = fsb.softmax_context_upsample(g["disp"].cuda(), g["spx"].cuda(), 4.0)
    assert tail.shape == g["up_tail"].shape
    assert rel_err(tail.cpu(), g["up_tail"]) <= TOL


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 104, 120, 160), (2, 48, 30, 41), (3, 7, 1, 5), (1, 1, 4, 4)])
def test_disparity_regression_vs_oracle(fsb, dtype, shape):
    B, D, H, W = shape
    g = torch.Generator().manual_seed(D * 31 + W)
    logits = (4.0 * torch.randn(B, D, H, W, generator=g)).to(dtype)
    prob = torch.softmax(logits.float(), 1).to(dtype)
    tol = {torch.float32: TOL, torch.float16: 2e-3, torch.bfloat16: 1.6e-2}[dtype]
    got = fsb.disparity_regression(prob.cuda(), D)
    want = oracle.disparity_regression(prob)
    assert got.dtype == dtype and got.shape == (B, 1, H, W)
    if D > 1:
        assert rel_err(got.float().cpu(), want.float()) <= tol
    else:
        assert got.abs().max().item() == 0
    got = fsb.softmax_disparity_regression(logits.cuda())
    want = oracle.softmax_disparity_regression(logits)
    assert got.dtype == torch.float32
    if D > 1:
        assert rel_err(got.cpu(), want) <= TOL  # the logits are exact in either dtype; the math is fp32
    else:
        assert got.abs().max().item() == 0


def test_softmax_regression_extremes(fsb):
    """-inf logits, a single finite logit per pixel, large magnitudes: no NaN, exact one-hot result."""
    B, D, H, W = 1, 40, 4, 8
    logits = torch.full((B, D, H, W), float("-inf"))
    idx = torch.arange(H * W).reshape(H, W) % D
    logits.scatter_(1, idx.reshape(1, 1, H, W), 0.0)
    got = fsb.softmax_disparity_regression(logits.cuda()).cpu()
    assert torch.equal(got[0, 0], idx.float())
    big = torch.zeros(B, D, H, W)
    big[:, 3] = 1e4
    big[:, 17] = -1e4
    assert torch.equal(fsb.softmax_disparity_regression(big.cuda()).cpu(), torch.full((B, 1, H, W), 3.0))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 120, 160), (2, 7, 9), (1, 1, 1), (3, 2, 33)])
def test_context_upsample_vs_oracle(fsb, dtype, shape):
    B, h, w = shape
    g = torch.Generator().manual_seed(h * 13 + w)
    disp = 100.0 * torch.rand(B, 1, h, w, generator=g)
    spx = (3.0 * torch.randn(B, 9, 4 * h, 4 * w, generator=g)).to(dtype)
    wts = torch.softmax(spx.float(), 1).to(dtype)
    tol = {torch.float32: TOL, torch.float16: 2e-3, torch.bfloat16: 1.6e-2}[dtype]
    got = fsb.context_upsample(disp.cuda(), wts.cuda())
    want = oracle.context_upsample(disp, wts)
    assert got.shape == (B, 4 * h, 4 * w) and got.dtype == torch.float32
    assert rel_err(got.cpu(), want.float()) <= tol
    got = fsb.softmax_context_upsample(disp.cuda(), spx.cuda(), 4.0)
    want = oracle.upsample_disp_tail(disp, spx)
    assert got.shape == (B, 1, 4 * h, 4 * w)
    assert rel_err(got.cpu(), want) <= TOL


def test_context_upsample_properties_full_size(fsb):
    """480x640: convex weights reproduce a constant field away from the border (weights sum to 1), a one-hot
    centre weight is 4x nearest-neighbour upsampling, and the op is linear in the disparity."""
    B, h, w = 2, 120, 160
    g = torch.Generator().manual_seed(3)
    spx = torch.randn(B, 9, 4 * h, 4 * w, generator=g).cuda()
    const = torch.full((B, 1, h, w), 7.5).cuda()
    up = fsb.softmax_context_upsample(const, spx, 4.0)
    assert (up[:, :, 4:-4, 4:-4] - 30.0).abs().max().item() <= 1e-4
    onehot = torch.zeros(B, 9, 4 * h, 4 * w)
    onehot[:, 4] = 1.0
    d = (50.0 * torch.rand(B, 1, h, w, generator=g)).cuda()
    nn = fsb.context_upsample(d, onehot.cuda())
    assert torch.equal(nn, d[:, 0].repeat_interleave(4, 1).repeat_interleave(4, 2))
    d2 = (50.0 * torch.rand(B, 1, h, w, generator=g)).cuda()
    wts = torch.softmax(spx, 1)
    lhs = fsb.context_upsample(d + d2, wts)
    rhs = fsb.context_upsample(d, wts) + fsb.context_upsample(d2, wts)
    assert rel_err(lhs.cpu(), rhs.cpu()) <= 1e-5


def test_install_rebinds_post_names_and_upsample(fsb):
    """patch.install also rebinds disparity_regression / context_upsample, and fused_upsample swaps
    model.upsample_disp (core/foundation_stereo.py:183-191) for the one-launch tail."""
    import types
    mod = types.ModuleType("core.foundation_stereo")
    mod.disparity_regression = mod.context_upsample = object()
    fsb.install(modules=[mod])
    assert mod.disparity_regression is fsb.disparity_regression and mod.context_upsample is fsb.context_upsample
    from foundationstereo_b200 import patch

    class Stub(torch.nn.Module):           # the two spx networks reduced to stand-ins with the right shapes
        def __init__(self):
            super().__init__()
            self.args = types.SimpleNamespace(mixed_precision=True)
            self.spx_gru = torch.nn.ConvTranspose2d(8, 9, 4, 2, 1)

        def spx_2_gru(self, mask, stem):
            return torch.nn.functional.interpolate(mask, scale_factor=2.0)[:, :8] + stem[:, :8]

        def upsample_disp(self, disp, mask, stem):
            with torch.autocast("cuda", enabled=True):
                p = torch.softmax(self.spx_gru(self.spx_2_gru(mask, stem)), 1)
                return oracle_free_upsample(disp * 4., p).unsqueeze(1).float()

    def oracle_free_upsample(disp, p):     # the reference's formula in torch ops on the GPU
        b, c, h, w = disp.shape
        u = torch.nn.functional.unfold(disp, 3, 1, 1).reshape(b, -1, h, w)
        u = torch.nn.functional.interpolate(u, (h * 4, w * 4), mode="nearest").reshape(b, 9, h * 4, w * 4)
        return (u * p).sum(1)

    m = Stub().cuda()
    disp = 40.0 * torch.rand(1, 1, 6, 10, device="cuda")
    mask = torch.randn(1, 32, 6, 10, device="cuda")
    stem = torch.randn(1, 32, 12, 20, device="cuda")
    with torch.no_grad():
        want = m.upsample_disp(disp, mask, stem)
        patch.install_upsample(m)
        got = m.upsample_disp(disp, mask, stem)
        patch.uninstall(m)
        again = m.upsample_disp(disp, mask, stem)
    assert got.shape == want.shape == (1, 1, 24, 40) and got.dtype == torch.float32
    assert rel_err(got.cpu(), want.cpu()) <= 1e-4
    assert torch.equal(again, want)
