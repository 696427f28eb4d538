import os, sys, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from foundationstereo_b200 import ops, _lib
from foundationstereo_b200._lib import FSB_PREC_TF32X3
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
C, H, W, L = 128, 120, 160, 2
g = torch.Generator().manual_seed(0)
f1 = torch.relu(torch.randn(B, C, H, W, generator=g)).cuda(); f2 = torch.relu(torch.randn(B, C, H, W, generator=g)).cuda()
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
levels = [torch.empty((B * H * W, W >> l), device="cuda") for l in range(L)]
for _ in range(3):
    ops.corr_pyramid_into(f1, f2, levels, FSB_PREC_TF32X3)
flush.zero_(); torch.cuda.synchronize()
ops.corr_pyramid_into(f1, f2, levels, FSB_PREC_TF32X3); torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * (64 * 16))()
L_ = _lib.lib(); L_.fsb_corr_trace_read.argtypes = [ctypes.c_void_p]; print("rc", L_.fsb_corr_trace_read(buf))
names = ["start", "setup done", "chunk0 arrived", "chunk0 lo done", "chunk1 arrived", "last lo done", "norms done", "acc ready", "stores issued"]
t00 = min(buf[b * 16] for b in range(32))
for b in (0, 1, 2, 15, 31):
    t = [buf[b * 16 + i] for i in range(9)]
    print(f"bh {b}: start +{(t[0]-t00)/1e3:.1f} us |", " ".join(f"{names[i]} {(t[i]-t[0])/1e3:.2f}" for i in range(1, 9)))
