"""Times the cost volume fused with corr_stem[0] against kernel 1 + torch.cat-free volume + the stock 1x1x1 Conv3d.
    python tools/volstem_exp.py [pairs]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import foundationstereo_b200 as fsb
from foundationstereo_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
C, H, W, D, G, Cp, Co = 128, 120, 160, 104, 8, 12, 28
g = torch.Generator(device="cuda").manual_seed(0)
for dt in (torch.float16, torch.float32):
    fl, fr = (torch.randn(B, C, H, W, device="cuda", generator=g).relu().to(dt) for _ in range(2))
    pl, pr = (torch.randn(B, Cp, H, W, device="cuda", generator=g).to(dt) for _ in range(2))
    wst = (torch.randn(Co, G + 2 * Cp, 1, 1, 1, device="cuda", generator=g) / 32 ** 0.5)
    bst = torch.randn(Co, device="cuda", generator=g) * 0.1

    def timeit(fn, reps=10):
        for _ in range(2):
            fn()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        torch.cuda.synchronize()
        ev[0].record()
        for _ in range(reps):
            fn()
        ev[1].record()
        torch.cuda.synchronize()
        return 1e3 * ev[0].elapsed_time(ev[1]) / reps

    def stock():
        v = torch.cat([fsb.build_gwc_volume(fl, fr, D, G), fsb.build_concat_volume(pl, pr, D)], 1)
        with torch.autocast("cuda", enabled=dt == torch.float16):
            return F.conv3d(v, wst, bst)

    def combined():
        v = ops.combined_volume(fl, fr, pl, pr, D, G)
        with torch.autocast("cuda", enabled=dt == torch.float16):
            return F.conv3d(v, wst, bst)

    fused = lambda: ops.volume_stem(fl, fr, pl, pr, wst, bst, D, G)
    a, b_, c = stock(), combined(), fused()
    err = float((c.float() - a.float()).abs().max() / a.float().abs().max())
    print(f"{str(dt):14s} pairs={B}: kernel 1 (two volumes) + cat + Conv3d {timeit(stock):8.1f} us | combined volume + Conv3d "
          f"{timeit(combined):8.1f} us | fused volume+stem {timeit(fused):8.1f} us ({c.numel() * c.element_size() / 1e6:.0f} MB out); "
          f"rel diff vs stock {err:.2e}")
