#!/usr/bin/env python
"""Per-launch device times of one update-block step (and the lookup) at a named shape:
CUDA events around every launch of the plan, L2 left warm as in the real loop.

    python tools/prof_update.py [H W [reps]]      (default 120 160 20)

Prints one line per launch: us, GFLOP, TFLOP/s; and the totals.  Also the short program that
`ncu --set full -k regex:conv_umma` is pointed at (reps=1)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import foundationstereo_b200 as fsb
from foundationstereo_b200 import synthetic

H = int(sys.argv[1]) if len(sys.argv) > 1 else 120
W = int(sys.argv[2]) if len(sys.argv) > 2 else 160
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
B = int(sys.argv[4]) if len(sys.argv) > 4 else 1
L, r, Cg = 2, 4, 28

state = synthetic.random_update_block_state(seed=0, corr_levels=L, corr_radius=r, volume_dim=Cg)
bn = {}
for kv in os.environ.get("FSB_BN", "").split(","):
    if "=" in kv:
        k, v = kv.split("=")
        bn[k] = int(v)
halo = os.environ.get("FSB_HALO")
halo = None if halo is None else ([] if halo == "none" else halo.split(","))
ub = fsb.SelectiveUpdateBlock(state, "cuda", use_graph=False, block_n=bn or None, halo=halo)
x = synthetic.hot_path_inputs(B, 16, H, W, 8)
c = lambda t: t.cuda()
net, inp, att = [c(t) for t in x["net"]], [c(t) for t in x["inp"]], [c(t) for t in x["att"]]
disp = c(x["disp"])
planes = L * (2 * r + 1) * (Cg + 1)
corr = torch.randn(B, planes, H, W, device="cuda")
for _ in range(2):
    net, mask, delta = ub(net, inp, corr, disp, att)
torch.cuda.synchronize()
plan = ub.plan_for(B, H, W)
steps = [(n_, f_) for n_, f_ in zip(plan.names, plan.steps) if callable(f_)]

n = len(steps)
tot = [0.0] * n
for _ in range(reps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    ev[0].record()
    for i, (_, fn) in enumerate(steps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    for i in range(n):
        tot[i] += ev[i].elapsed_time(ev[i + 1])
rows = []
sum_us, sum_gf = 0.0, 0.0
for i, (name, fn) in enumerate(steps):
    us = 1e3 * tot[i] / reps
    gf = getattr(fn, "flops", 0.0) / 1e9
    sum_us += us
    sum_gf += gf
    rows.append({"name": name, "us": round(us, 2), "gflop": round(gf, 3), "tflops": round(gf / us * 1e3, 1) if us > 0 else 0})
    print(f"{name:32s} {us:9.2f} us  {gf:8.3f} GF  {gf / us * 1e3:8.1f} TFLOP/s" if gf else f"{name:32s} {us:9.2f} us")
print(f"{'TOTAL (event-separated)':32s} {sum_us:9.2f} us  {sum_gf:8.3f} GF  {sum_gf / sum_us * 1e3:8.1f} TFLOP/s")

# back-to-back (no events between launches) and graph replay
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    ub.run_steps(plan)
e1.record()
torch.cuda.synchronize()
b2b = 1e3 * e0.elapsed_time(e1) / reps
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    ub.run_steps(plan)
g.replay()
torch.cuda.synchronize()
e0.record()
for _ in range(reps):
    g.replay()
e1.record()
torch.cuda.synchronize()
gr = 1e3 * e0.elapsed_time(e1) / reps
print(f"back-to-back {b2b:.1f} us/step, graph replay {gr:.1f} us/step ({sum_gf / gr * 1e3:.1f} TFLOP/s)")
out = {"shape": [B, H, W], "launches": rows, "sum_us": sum_us, "gflop": sum_gf, "back_to_back_us": b2b, "graph_us": gr}
os.makedirs("gpurun_out", exist_ok=True)
with open(f"gpurun_out/prof_update_{H}x{W}.json", "w") as fh:
    json.dump(out, fh, indent=1)
