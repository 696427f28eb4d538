"""GPU tests of kernel 4 (tcgen05 implicit-GEMM convolution, csrc/conv.cu) and the NHWC
helpers (csrc/nhwc.cu) through the C ABI, against plain PyTorch fp32 references of the
same ops evaluated on the SAME 16-bit-rounded operands.

Tolerance: products of 16-bit operands are exact in fp32, so the only differences are the
accumulation order (fp32) and the final rounding of a 16-bit output: 2^-10 relative for
fp16 outputs, 2^-7 for bf16; fp32 outputs agree to 1e-4 relative.
"""
import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fsb():
    import foundationstereo_b200 as f
    return f


def _tol(dtype):
    return 2e-3 if dtype == torch.float16 else 1.6e-2


def _nhwc(x, cpad, dtype):
    """(B,C,H,W) fp32 -> (B,H,W,cpad) 16-bit on the GPU (pad channels zero)."""
    B, Cc, H, W = x.shape
    out = torch.zeros(B, H, W, cpad, dtype=dtype, device="cuda")
    out[..., :Cc] = x.permute(0, 2, 3, 1).to(dtype)
    return out


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g) * scale


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("cfg", [
    # B, H, W, [source channels], Cout, ksize, block_n, epilogue
    (1, 16, 24, [64], 64, 3, 64, "relu"),
    (2, 9, 13, [128, 64], 127, 3, 64, "relu"),       # ragged tiles, Cout not a multiple of 32
    (1, 30, 40, [128, 128, 128], 384, 3, 128, "relu"),
    (1, 8, 12, [576], 256, 1, 128, "relu"),            # 1x1 (convc1 shape)
    (1, 20, 20, [128], 512, 1, 256, "gelu"),
    (1, 17, 31, [64], 32, 3, 32, "bias"),
    (1, 5, 7, [128], 1, 3, 16, "bias"),                # disparity head output
])
@pytest.mark.parametrize("layout", [0, 1])  # FSB_CONV_LAYOUT_TAP / FSB_CONV_LAYOUT_HALO
def test_conv_vs_torch(fsb, dtype, cfg, layout):
    from foundationstereo_b200 import _lib
    from foundationstereo_b200.conv import ConvLaunch, pack_rows, pack_segment
    B, H, W, cins, cout, k, bn, epi = cfg
    xs = [_rand(B, c, H, W, seed=10 + i).to(dtype).float() for i, c in enumerate(cins)]
    cin = sum(cins)
    w = (_rand(cout, cin, k, k, seed=3) / (cin * k * k) ** 0.5).to(dtype).float()
    b = _rand(cout, seed=4)
    want = F.conv2d(torch.cat(xs, 1).double(), w.double(), b.double(), padding=k // 2).float()
    want = {"relu": F.relu, "gelu": F.gelu, "bias": lambda t: t}[epi](want)
    srcs = [_nhwc(x, x.shape[1], dtype) for x in xs]
    rows = (cout + bn - 1) // bn * bn
    wp = pack_rows([pack_segment(w, [(c, c) for c in cins], layout)], rows).to(dtype).cuda().contiguous()
    bp = torch.zeros(1, rows)
    bp[0, :cout] = b
    ld = (cout + 7) // 8 * 8 + 8
    out = torch.full((B, H, W, ld), 7.0, dtype=dtype, device="cuda")
    out32 = torch.empty(B, cout, H, W, device="cuda")
    code = {"relu": _lib.FSB_EPI_RELU, "gelu": _lib.FSB_EPI_GELU, "bias": _lib.FSB_EPI_BIAS}[epi]
    n0 = fsb.launch_count()
    ConvLaunch(srcs=srcs, segs=[(k, list(range(len(cins))))], weight=wp, bias=bp.cuda(), cout=cout, block_n=bn,
               epilogue=code, out=out, ch_off=8, out_f32=out32, out_scale=0.5, layout=layout)()
    torch.cuda.synchronize()
    assert fsb.launch_count() == n0 + 1
    assert rel_err(out32.cpu(), 0.5 * want) <= 1e-4
    got16 = out[..., 8:8 + cout].permute(0, 3, 1, 2).float().cpu()
    assert rel_err(got16, want) <= _tol(dtype)
    # channels outside [ch_off, ch_off + Cout) are untouched
    assert (out[..., :8] == 7.0).all() and (out[..., 8 + cout:] == 7.0).all()


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_conv_scale_residual(fsb, dtype):
    from foundationstereo_b200 import _lib
    from foundationstereo_b200.conv import ConvLaunch, pack_rows, pack_segment
    B, H, W, cin, cout = 1, 11, 19, 512, 128
    x = _rand(B, cin, H, W, seed=1).to(dtype).float()
    res = _rand(B, cout, H, W, seed=2).to(dtype).float()
    w = (_rand(cout, cin, seed=3) / cin ** 0.5).to(dtype).float()
    b, gamma = _rand(cout, seed=4), _rand(cout, seed=5)
    want = res + gamma[None, :, None, None] * F.conv2d(x, w[:, :, None, None], b)
    out = torch.empty(B, H, W, cout, dtype=dtype, device="cuda")
    ConvLaunch(srcs=[_nhwc(x, cin, dtype)], segs=[(1, [0])],
               weight=pack_rows([pack_segment(w, [(cin, cin)])], cout).to(dtype).cuda(), bias=b[None].cuda(), cout=cout,
               block_n=64, epilogue=_lib.FSB_EPI_SCALE_RES, out=out, res=_nhwc(res, cout, dtype), gamma=gamma.cuda())()
    assert rel_err(out.permute(0, 3, 1, 2).float().cpu(), want) <= _tol(dtype)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 30, 40, 256), (2, 7, 9, 384), (1, 40, 20, 384)])
@pytest.mark.parametrize("layout", [0, 1])
def test_gru_gates_vs_torch(fsb, dtype, shape, layout):
    """The two gate launches reproduce RaftConvGRU small (1x1) + large (3x3) and the
    SelectiveConvGRU blend (core/update.py:90-95,119) from the same hx / x / h."""
    from foundationstereo_b200 import _lib
    from foundationstereo_b200.conv import ConvLaunch, interleave_zr, pack_rows, pack_segment
    B, H, W, cin = shape
    hd = 128
    q16 = lambda t: t.to(dtype).float()
    h = torch.tanh(_rand(B, hd, H, W, seed=1))
    x = q16(torch.relu(_rand(B, cin, H, W, seed=2)))
    hx = q16(torch.relu(_rand(B, cin + hd, H, W, seed=3)))
    att = torch.sigmoid(_rand(B, 1, H, W, seed=4))
    W_ = {}
    for gi, (gru, k) in enumerate((("l", 3), ("s", 1))):
        for gj, gate in enumerate("zrq"):
            W_[gru + gate] = (q16(_rand(hd, cin + hd, k, k, seed=20 + 3 * gi + gj) / ((cin + hd) * k * k) ** 0.5),
                              _rand(hd, seed=40 + 3 * gi + gj) * 0.1)

    def raft(gru, k):
        z = torch.sigmoid(F.conv2d(hx, *W_[gru + "z"], padding=k // 2))
        r = torch.sigmoid(F.conv2d(hx, *W_[gru + "r"], padding=k // 2))
        rh = q16(r * h)  # the kernel stores r*h as a 16-bit operand of the q convolution
        q = torch.tanh(F.conv2d(torch.cat([rh, x], 1), *W_[gru + "q"], padding=k // 2))
        zq = q16(z)      # z is kept 16-bit between the two launches
        return (1 - zq) * h + zq * q
    want = raft("s", 1) * att + raft("l", 3) * (1 - att)

    bn = 128
    new = lambda c, dt=dtype: torch.empty(B, H, W, c, dtype=dt, device="cuda")
    z_l, z_s, rh_l, rh_s = new(hd), new(hd), new(hd), new(hd)
    h32 = h.permute(0, 2, 3, 1).contiguous().cuda()
    hx16, x16 = _nhwc(hx, cin + hd, dtype), _nhwc(x, cin, dtype)
    hxs = [(cin + hd, cin + hd)]
    zr_w = [interleave_zr(pack_segment(W_[g + "z"][0], hxs, layout), pack_segment(W_[g + "r"][0], hxs, layout), bn)
            for g in "ls"]
    zr_b = torch.stack([interleave_zr(W_[g + "z"][1][:, None], W_[g + "r"][1][:, None], bn).flatten() for g in "ls"])
    ConvLaunch(srcs=[hx16], segs=[(3, [0]), (1, [0])], weight=pack_rows(zr_w, 2 * hd).to(dtype).cuda(), bias=zr_b.cuda(),
               cout=2 * hd, block_n=bn, epilogue=_lib.FSB_EPI_GATE_ZR, h32=h32, z_l=z_l, z_s=z_s, rh_l=rh_l, rh_s=rh_s,
               hidden=hd, layout=layout)()
    rhx = [(hd, hd), (cin, cin)]
    q_w = [pack_segment(W_[g + "q"][0], rhx, layout) for g in "ls"]
    q_b = torch.stack([W_[g + "q"][1] for g in "ls"])
    h16 = new(hd)
    h32_out = torch.empty_like(h32)
    ConvLaunch(srcs=[rh_l, rh_s, x16], segs=[(3, [0, 2]), (1, [1, 2])], weight=pack_rows(q_w, hd).to(dtype).cuda(),
               bias=q_b.cuda(), cout=hd, block_n=64, epilogue=_lib.FSB_EPI_GATE_Q, h32=h32, att=att.reshape(B, H, W).cuda(),
               z_l=z_l, z_s=z_s, h32_out=h32_out, h16_out=h16, hidden=hd, layout=layout)()
    got = h32_out.permute(0, 3, 1, 2).cpu()
    # r*h and z are rounded to 16 bits at slightly different values than in the torch restatement
    # (fp32 accumulation order), so allow a few output ulps
    assert rel_err(got, want) <= 3 * _tol(dtype)
    assert rel_err(h16.permute(0, 3, 1, 2).float().cpu(), want) <= 4 * _tol(dtype)


# ------------------------------------------------------------------------------- helpers
def test_layout_helpers(fsb):
    from foundationstereo_b200 import conv as C
    x = _rand(2, 37, 9, 11, seed=1).cuda()
    d16 = torch.full((2, 9, 11, 64), 5.0, dtype=torch.float16, device="cuda")
    d32 = torch.empty(2, 9, 11, 37, device="cuda")
    C.nchw_to_nhwc(x, d16, d32)
    assert torch.equal(d32, x.permute(0, 2, 3, 1))
    assert torch.equal(d16[..., :37], x.permute(0, 2, 3, 1).half())
    assert (d16[..., 37:] == 0).all()
    back = torch.empty_like(x)
    C.nhwc_to_nchw(d32, back)
    assert torch.equal(back, x)


@pytest.mark.parametrize("hw", [(30, 40), (15, 21), (8, 12)])
def test_pool_and_resize(fsb, hw):
    from foundationstereo_b200 import conv as C
    H, W = hw
    x = _rand(2, 128, H, W, seed=2).cuda()
    x_nhwc = x.permute(0, 2, 3, 1).contiguous()
    Ho, Wo = (H + 1) // 2, (W + 1) // 2
    p = torch.empty(2, Ho, Wo, 128, dtype=torch.float16, device="cuda")
    C.pool2x(x_nhwc, p)
    want = F.avg_pool2d(x, 3, stride=2, padding=1)
    assert want.shape[2:] == (Ho, Wo)
    assert rel_err(p.permute(0, 3, 1, 2).float(), want) <= 1e-3
    small = want.permute(0, 2, 3, 1).contiguous()
    up = torch.empty(2, H, W, 128, dtype=torch.float16, device="cuda")
    C.resize(small, up)
    want_up = F.interpolate(want, (H, W), mode="bilinear", align_corners=True)
    assert rel_err(up.permute(0, 3, 1, 2).float(), want_up) <= 1e-3
    # 16-bit sources (what the engine reads: the fp16 / bf16 copy of the hidden state) against fp32 PyTorch on the same
    # rounded values
    for dt, tol in ((torch.float16, 1e-3), (torch.bfloat16, 8e-3)):
        x16 = x_nhwc.to(dt)
        p16 = torch.empty(2, Ho, Wo, 128, dtype=dt, device="cuda")
        C.pool2x(x16, p16)
        want16 = F.avg_pool2d(x16.float().permute(0, 3, 1, 2), 3, stride=2, padding=1)
        assert rel_err(p16.permute(0, 3, 1, 2).float(), want16) <= tol
        s16 = want16.permute(0, 2, 3, 1).contiguous().to(dt)
        up16 = torch.empty(2, H, W, 128, dtype=dt, device="cuda")
        C.resize(s16, up16)
        want_up16 = F.interpolate(s16.float().permute(0, 3, 1, 2), (H, W), mode="bilinear", align_corners=True)
        assert rel_err(up16.permute(0, 3, 1, 2).float(), want_up16) <= tol


def test_conv7x7_helpers(fsb):
    from foundationstereo_b200 import conv as C
    B, H, W = 2, 13, 37
    disp = (_rand(B, 1, H, W, seed=1) * 20).cuda()
    w, b = (_rand(64, 1, 7, 7, seed=2) / 7).cuda(), _rand(64, seed=3).cuda()
    out = torch.empty(B, H, W, 64, dtype=torch.float16, device="cuda")
    motion = torch.zeros(B, H, W, 128, dtype=torch.float16, device="cuda")
    C.conv7x7_c1(disp, w.reshape(64, 49).contiguous(), b, out, 64, motion, 127)
    want = F.relu(F.conv2d(disp, w, b, padding=3))
    assert rel_err(out.permute(0, 3, 1, 2).float(), want) <= 1e-3
    assert torch.equal(motion[..., 127], disp[:, 0].half()) and (motion[..., :127] == 0).all()
    x = _rand(B, 128, H, W, seed=4).half().cuda()
    wd, bd = (_rand(128, 1, 7, 7, seed=5) / 7).cuda(), _rand(128, seed=6).cuda()
    o = torch.empty(B, H, W, 128, dtype=torch.float16, device="cuda")
    C.dwconv7x7(x.permute(0, 2, 3, 1).contiguous(), wd.reshape(128, 49).contiguous(), bd, o)
    want = F.conv2d(x.float(), wd, bd, padding=3, groups=128)
    assert rel_err(o.permute(0, 3, 1, 2).float(), want) <= 1e-3
    y = _rand(1000, seed=7).cuda()
    y0 = y.clone()
    xx = _rand(1000, seed=8).cuda()
    C.axpy(y, xx)
    assert torch.equal(y, y0 + xx)


# ------------------------------------------------------------------ fused EdgeNeXt MLP (csrc/mlp.cu)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("npix", [128, 300, 148 * 128 * 3 + 77])  # one tile, ragged, several tiles per persistent CTA
def test_edgenext_mlp_vs_torch(fsb, dtype, npix):
    """out = res + gamma * (GELU(x W1^T + b1) W2^T + b2) (core/submodule.py:584-590) against fp32 PyTorch on
    the same 16-bit-rounded operands; the hidden activations are rounded to the 16-bit type in between, as the
    reference's autocast Linear does."""
    from foundationstereo_b200 import conv as Cv
    C_, F_ = 128, 512
    x = _rand(npix, C_, seed=1).to(dtype)
    res = _rand(npix, C_, seed=2).to(dtype)
    w1 = (_rand(F_, C_, seed=3) / C_ ** 0.5).to(dtype)
    w2 = (_rand(C_, F_, seed=4) / F_ ** 0.5).to(dtype)
    b1, b2, gamma = _rand(F_, seed=5) * 0.1, _rand(C_, seed=6) * 0.1, _rand(C_, seed=7)
    h = F.gelu(x.float() @ w1.float().t() + b1).to(dtype).float()
    want = res.float() + gamma * (h @ w2.float().t() + b2)
    out = torch.empty(npix, C_, dtype=dtype, device="cuda")
    c = lambda t: t.cuda()
    Cv.edgenext_mlp(c(x), c(w1), c(b1), c(w2), c(b2), c(gamma), c(res), out)
    torch.cuda.synchronize()
    assert rel_err(out.float().cpu(), want) <= 2 * _tol(dtype)


def test_update_block_fused_mlp_equals_two_launches(fsb):
    """The engine with the fused MLP gives the same step as with pwconv1 / pwconv2 as separate convolutions (same
    operands, same roundings: only the fp32 accumulation order inside the GEMMs may differ)."""
    import oracle
    p = oracle.make_update_block_params(seed=3)
    x = oracle.make_inputs(B=1, C=16, H=24, W=40, D=8, Cg=28, seed=9)
    c = lambda t: t.cuda()
    outs = []
    for fused in (True, False):
        ub = fsb.SelectiveUpdateBlock(p, "cuda", use_graph=False, fused_mlp=fused)
        corr = _rand(1, ub.planes, 24, 40, seed=5).cuda()
        net, mask, delta = ub([c(t) for t in x["net"]], [c(t) for t in x["inp"]], corr, c(x["disp"]), [c(t) for t in x["att"]])
        outs.append((delta.clone(), mask.clone()))
    assert rel_err(outs[0][0].cpu(), outs[1][0].cpu()) <= 4e-3
    assert torch.equal(outs[0][1], outs[1][1])


# ------------------------------------------------------------------ DispHead.conv[4] on CUDA cores (csrc/nhwc.cu)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(2, 13, 37, 128), (1, 8, 32, 128), (3, 17, 70, 64), (1, 1, 1, 32)])
def test_conv3x3_cout1_vs_torch(fsb, dtype, shape):
    """3x3 convolution C -> 1 (core/update.py:28) against fp32 PyTorch on the same 16-bit-rounded operands: fp32
    accumulation, fp32 output -> 1e-4 relative; with disp_accum the same launch adds delta to the disparity."""
    from foundationstereo_b200 import conv as C
    B, H, W, Cc = shape
    x = _rand(B, Cc, H, W, seed=1).to(dtype)
    w = (_rand(1, Cc, 3, 3, seed=2) / (3 * Cc ** 0.5)).to(dtype)
    b = _rand(1, seed=3)
    want = F.conv2d(x.float(), w.float(), b, padding=1)
    wt = w[0].float().permute(1, 2, 0).reshape(9, Cc).contiguous().cuda()
    xd = x.permute(0, 2, 3, 1).contiguous().cuda()
    delta = torch.full((B, 1, H, W), 7.0, device="cuda")
    C.conv3x3_cout1(xd, wt, b.cuda(), delta, None)
    assert rel_err(delta.cpu(), want) <= 1e-4
    disp = _rand(B, 1, H, W, seed=4).cuda()
    d0 = disp.clone()
    delta2 = torch.empty_like(delta)
    C.conv3x3_cout1(xd, wt, b.cuda(), delta2, disp)
    assert torch.equal(delta2, delta) and torch.equal(disp, d0 + delta)


def test_update_block_dot_head4_equals_tensor_core_head4(fsb):
    """disp_head.conv.4 as the CUDA-core dot kernel vs as a tensor-core convolution launch: same operands, fp32
    accumulation in both -> delta agrees to accumulation-order noise."""
    import oracle
    p = oracle.make_update_block_params(seed=4)
    x = oracle.make_inputs(B=2, C=16, H=20, W=44, D=8, Cg=28, seed=10)
    c = lambda t: t.cuda()
    outs = []
    for dot in (True, False):
        ub = fsb.SelectiveUpdateBlock(p, "cuda", use_graph=False, dot_head4=dot)
        corr = _rand(2, ub.planes, 20, 44, seed=5).cuda()
        net, mask, delta = ub([c(t) for t in x["net"]], [c(t) for t in x["inp"]], corr, c(x["disp"]), [c(t) for t in x["att"]])
        outs.append(delta.clone())
    assert rel_err(outs[0].cpu(), outs[1].cpu()) <= 1e-4


# ------------------------------------------------------------------ whole EdgeNeXt block in one launch (csrc/edgenext.cu)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 8, 16), (2, 13, 37), (1, 24, 40), (3, 120, 160)])  # one tile, ragged, several tiles per CTA
def test_edgenext_block_equals_dwconv_plus_mlp(fsb, dtype, shape):
    """fsb_edgenext_block == fsb_dwconv7x7_nhwc followed by fsb_edgenext_mlp, BIT for bit (same accumulation order in
    the depthwise part, the same 16-bit operand tile feeds the same MMA sequence), and against fp32 PyTorch."""
    from foundationstereo_b200 import conv as Cv
    B, H, W = shape
    C_, F_ = 128, 512
    x = _rand(B, H, W, C_, seed=1).to(dtype).cuda()
    dw_w, dw_b = (_rand(C_, 49, seed=2) / 7).cuda(), (_rand(C_, seed=3) * 0.1).cuda()
    w1 = (_rand(F_, C_, seed=4) / C_ ** 0.5).to(dtype).cuda()
    w2 = (_rand(C_, F_, seed=5) / F_ ** 0.5).to(dtype).cuda()
    b1, b2, gamma = (_rand(F_, seed=6) * 0.1).cuda(), (_rand(C_, seed=7) * 0.1).cuda(), _rand(C_, seed=8).cuda()
    t = torch.empty_like(x)
    Cv.dwconv7x7(x, dw_w, dw_b, t)
    two = torch.empty_like(x)
    Cv.edgenext_mlp(t, w1, b1, w2, b2, gamma, x, two)
    one = torch.full_like(x, 7.0)
    Cv.edgenext_block(x, dw_w, dw_b, w1, b1, w2, b2, gamma, one)
    torch.cuda.synchronize()
    assert torch.equal(one, two), float((one.float() - two.float()).abs().max())
    xt = x.float().permute(0, 3, 1, 2)
    tt = F.conv2d(xt, dw_w.reshape(C_, 1, 7, 7), dw_b, padding=3, groups=C_).permute(0, 2, 3, 1).to(dtype).float()
    h = F.gelu(tt @ w1.float().t() + b1).to(dtype).float()
    want = x.float() + gamma * (h @ w2.float().t() + b2)
    assert rel_err(one.float().cpu(), want.cpu()) <= 2 * _tol(dtype)
