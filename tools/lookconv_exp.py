"""Times the lookup, encoder.convc1 and the fused lookup -> convc1 launch alone (B pairs of 120x160, D=104, Cg=28, L=2, r=4).
    python tools/lookconv_exp.py [pairs]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import foundationstereo_b200 as fsb
from foundationstereo_b200 import ops, conv as Cv, _lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
H, W, D, Cg, L, r, K = 120, 160, 104, 28, 2, 4, 9
dt = torch.float16
g = torch.Generator(device="cuda").manual_seed(0)
npix = B * H * W
geo_levels = [torch.randn(npix, D >> l, Cg, device="cuda", generator=g) for l in range(L)]
corr_levels = [torch.randn(npix, W >> l, device="cuda", generator=g) for l in range(L)]
disp = torch.rand(B, 1, H, W, device="cuda", generator=g) * D
coords = torch.arange(W, dtype=torch.float32, device="cuda").reshape(1, 1, W, 1).repeat(B, H, 1, 1)
P = L * K * (Cg + 1)
cpad = (P + 63) // 64 * 64
wp = torch.zeros(256, cpad, dtype=dt, device="cuda")
wp[:, :P] = (torch.randn(256, P, device="cuda", generator=g) / P ** 0.5).to(dt)
bias = torch.randn(256, device="cuda", generator=g) * 0.1
wl = ops.pack_convc1_weight(wp[:, :P].float(), L, K, Cg, dt)
buf = torch.empty(B, H, W, cpad, dtype=dt, device="cuda")
out1, out2 = torch.empty(B, H, W, 256, dtype=dt, device="cuda"), torch.empty(B, H, W, 256, dtype=dt, device="cuda")
convs = {bn: Cv.ConvLaunch(srcs=[buf], segs=[(1, [0])], weight=wp, bias=bias, cout=256, block_n=bn, epilogue=_lib.FSB_EPI_RELU,
                           name="convc1", layout=_lib.FSB_CONV_LAYOUT_TAP, out=out1) for bn in (128, 256)}


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(reps):
        fn()
    ev[1].record()
    torch.cuda.synchronize()
    return 1e3 * ev[0].elapsed_time(ev[1]) / reps


look = lambda: ops.geo_lookup_nhwc16(geo_levels, corr_levels, disp, coords, r, buf)
fused = lambda: ops.lookup_convc1(geo_levels, corr_levels, disp, coords, r, wl, bias, out2)
t_look = timeit(look)
t_conv = {bn: timeit(c) for bn, c in convs.items()}
t_fused = timeit(fused)
look(); convs[128](); fused(); torch.cuda.synchronize()
err = float((out1.float() - out2.float()).abs().max() / out1.float().abs().max())
rd = sum(t.numel() for t in geo_levels) * 0  # (window bytes below)
win = npix * (4 + L * (Cg + 1) * (K + 1) * 4)
print(f"pairs={B}: lookup {t_look:.1f} us ({(win + npix * cpad * 2) / t_look * 1e-3:.0f} GB/s), convc1 bn128 {t_conv[128]:.1f} us, "
      f"bn256 {t_conv[256]:.1f} us, fused {t_fused:.1f} us ({(win + npix * 512) / t_fused * 1e-3:.0f} GB/s algorithmic), "
      f"rel diff fused vs two launches {err:.2e}, equal={torch.equal(out1, out2)}")
