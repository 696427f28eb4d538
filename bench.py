#!/usr/bin/env python
"""bench.py -- FoundationStereo hot path on B200: pairs/s @ 480x640, max_disp 416, 32 GRU iterations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one pass of the hot path over one synthetic stereo pair per GPU
(BASELINE.json configs[1] shapes: 1/4-res features 120x160, C=128, D=104, corr_levels 2,
radius 4, 32 iterations of lookup + update block, random-init weights):
    volumes (kernel 1) -> geometry-encoding volume (kernels 2, 2b) -> 32 x (kernel 3 + update block).
The backbone and the 3-D hourglass of the reference are out of scope (BASELINE.json), so the
tensors they would produce (features, filtered volume, context) are seeded synthetic inputs.

One JSON line on stdout (rank 0).  `value` = pairs/s with inputs resident in HBM;
`e2e` = the same through the public API with pinned host buffers, H2D + D2H inside the
timed region; `roofline` = the dominant kernel of libfsb200.so against MEASURED_PEAKS.json;
`cpu_baseline` = the CPU oracle (a port of the reference's algorithm) on the host cores.
`--impl reference` times that CPU path alone (rank 0 only).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CFG = dict(B=1, C=128, H=120, W=160, D=104, Cp=12, Cg=28, G=8, L=2, r=4, iters=32)
METRIC = "pairs/sec @480x640 D416 32it"
UNIT = "pairs/s"


def workload_name():
    return ("hot path of FoundationStereo ViT-S config (BASELINE configs[1]): 1/4-res 120x160, C=128, max_disp=416 "
            "(D=104), corr_levels=2, radius=4, 32 x (lookup + update block); a step = one batch of pairs per GPU "
            "(config.pairs_per_gpu_per_step)")


# --------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores (bounded sample, extrapolated linearly)
# --------------------------------------------------------------------------------------
def cpu_sample(rows: int, iters: int, threads: int):
    """Times the CPU oracle on a horizontal band of `rows` of the 120 image rows and `iters`
    of the 32 iterations; every stage of the path is linear in rows and in iterations, so
    pair time = (120/rows) * [t_volumes + t_geo + (32/iters) * t_iters]."""
    import torch
    import oracle
    torch.set_num_threads(threads)
    c = CFG
    x = oracle.make_inputs(B=1, C=c["C"], H=rows, W=c["W"], D=c["D"], Cp=c["Cp"], Cg=c["Cg"], seed=0)
    p = oracle.make_update_block_params(seed=0, corr_levels=c["L"], corr_radius=c["r"], volume_dim=c["Cg"],
                                        delta_scale=0.25)
    with torch.no_grad():
        t0 = time.perf_counter()
        oracle.gwc_volume(x["fl"], x["fr"], c["D"], c["G"])
        oracle.concat_volume(x["pl"], x["pr"], c["D"])
        t1 = time.perf_counter()
        geo = oracle.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], c["L"], c["r"])
        t2 = time.perf_counter()
        oracle.refine(p, geo, x["net"], x["inp"], x["att"], x["disp"], iters)
        t3 = time.perf_counter()
    scale = c["H"] / rows
    pair_s = scale * ((t1 - t0) + (t2 - t1) + (c["iters"] / iters) * (t3 - t2))
    return pair_s, (t3 - t0)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 1
    rows, iters = 30, 2
    for _ in range(args.warmup):
        cpu_sample(rows, iters, threads)
    pair_times, walls = [], []
    for _ in range(args.steps):
        pt, wall = cpu_sample(rows, iters, threads)
        pair_times.append(pt)
        walls.append(wall)
    pair_s = statistics.mean(pair_times)
    value = 1.0 / pair_s
    sample = (f"per step: rows 0..{rows - 1} of 120 (full width, D=104) through volumes + geo init + {iters} of 32 "
              f"iterations, extrapolated linearly to one pair; torch fp32, {threads} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * pair_s, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(), "l2": "cpu run"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU oracle (oracle/hotpath.py), a restatement of the reference's PyTorch code pinned to the "
                "reference's outputs; the reference itself is Python and /root/reference does not exist on the box",
    }
    emit(line)
    return 0


# --------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# --------------------------------------------------------------------------------------
class Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.rows.append((time.perf_counter(), line.strip()))
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, t0: float, t1: float):
        sm, smax, reasons = [], [], set()
        for t, line in self.rows:
            if not (t0 <= t <= t1 + 0.15):
                continue
            f = [s.strip() for s in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(smax), "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------
# algorithmic work per launch
# --------------------------------------------------------------------------------------
def algorithmic_bytes(c):
    B, C, H, W, D, Cp, Cg, G, L, r = (c[k] for k in ("B", "C", "H", "W", "D", "Cp", "Cg", "G", "L", "r"))
    K = 2 * r + 1
    px = B * H * W
    return {
        "volume_gwc_concat": px * ((2 * C + 2 * Cp) * 4 + (G + 2 * Cp) * D * 4),
        "corr_pyramid": 8 * px * C + 4 * px * sum(W >> l for l in range(L)),
        "geo_pyramid": 4 * px * Cg * (D + sum(D >> l for l in range(L))),
        "geo_lookup": px * (4 + L * (Cg + 1) * (K + 1) * 4 + L * K * (Cg + 1) * 4),
    }


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return {"hbm_gbs": p["hbm_gbs"], "bf16_tflops": p["bf16_tflops"],
                "bf16_tflops_sustained": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


# --------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------
def time_kernel(torch, fn, reps, flush):
    """average device time of one launch in ms, L2 flushed (a 512 MiB write) before each."""
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for _ in range(3):
        fn()
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / reps


def run_b200(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # convenience: `python bench.py --gpus N` re-launches itself one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__),
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup), "--pairs", str(args.pairs)]
        return subprocess.call(cmd)

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        emit({"error": "no CUDA device: the hot path has no CPU fallback"})
        return 2
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    import foundationstereo_b200 as fsb
    from foundationstereo_b200 import ops, synthetic
    from foundationstereo_b200._lib import FSB_PREC_FP32

    c = dict(CFG)
    c["B"] = args.pairs
    x = synthetic.hot_path_inputs(c["B"], c["C"], c["H"], c["W"], c["D"], c["Cp"], c["Cg"], seed=rank)
    state = synthetic.random_update_block_state(seed=0, corr_levels=c["L"], corr_radius=c["r"], volume_dim=c["Cg"])
    hp = fsb.StereoHotPath(state, max_disp=4 * c["D"], corr_levels=c["L"], corr_radius=c["r"], num_groups=c["G"],
                           device=device)

    def pin(t):
        return t.contiguous().pin_memory()
    host = {k: ([pin(t) for t in v] if isinstance(v, list) else pin(v)) for k, v in x.items()}

    def to_dev(h):
        return {k: ([t.to(device, non_blocking=True) for t in v] if isinstance(v, list)
                    else v.to(device, non_blocking=True)) for k, v in h.items()}
    dev = to_dev(host)
    h2d_bytes = sum(sum(t.numel() * t.element_size() for t in v) if isinstance(v, list)
                    else v.numel() * v.element_size() for v in host.values())

    def step(d):
        disp, mask, comb = hp.run(d["fl"], d["fr"], d["pl"], d["pr"], d["volume"], d["disp"], d["net"], d["inp"],
                                  d["att"], c["iters"])
        return disp

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident-input timing ---------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        out = step(dev)
    barrier()
    clocks = Clocks(local_rank)
    clocks.start()
    time.sleep(0.3)
    n0 = fsb.launch_count() + hp.update_block.replayed_kernels
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    tw0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        out = step(dev)
    e1.record()
    barrier()
    tw1 = time.perf_counter()
    launches = fsb.launch_count() + hp.update_block.replayed_kernels - n0  # eager launches + kernels inside graph replays
    ms_total = e0.elapsed_time(e1)

    # ---- end to end: pinned host buffers in, disparity out, every step -------------------
    # (StereoHotPath.run_pinned: the upload of step i+1 runs on a copy stream while step i computes; every step's
    # inputs cross PCIe inside the timed region and every step's disparity comes back)
    out_hosts = [torch.empty_like(out, device="cpu").pin_memory() for _ in range(max(args.steps, 2))]
    hp.run_pinned([host] * 2, c["iters"], out_hosts[:2])
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    te0 = time.perf_counter()
    f0.record()
    hp.run_pinned([host] * args.steps, c["iters"], out_hosts[:args.steps])
    f1.record()
    barrier()
    te1 = time.perf_counter()
    out_host = out_hosts[-1]
    clocks.stop()
    ms_e2e = max(f0.elapsed_time(f1), 1e3 * (te1 - te0))
    d2h_bytes = out_host.numel() * out_host.element_size()

    # whole-job figures: SUM of pairs, MAX of device times over ranks (foundationstereo_b200/shard.py)
    from foundationstereo_b200 import shard
    pairs, ms_total, value = shard.reduce_throughput(c["B"] * args.steps, ms_total, device)
    _, ms_e2e, e2e_value = shard.reduce_throughput(c["B"] * args.steps, ms_e2e, device)

    line = None
    if rank == 0:
        # ---- per-kernel device times for the roofline (rank 0, after the timed regions) ------
        peaks = measured_peaks()
        flush = torch.empty(512 << 20, dtype=torch.uint8, device=device)
        D, G, L, r = c["D"], c["G"], c["L"], c["r"]
        geo = fsb.Combined_Geo_Encoding_Volume(dev["fl"], dev["fr"], dev["volume"], num_levels=L, dx=hp.dx)
        coords = torch.arange(c["W"], dtype=torch.float32, device=device).reshape(1, 1, -1, 1).repeat(c["B"], c["H"], 1, 1)
        from foundationstereo_b200._lib import FSB_PREC_TF32X3
        plan_b = hp.update_block.plan_for(c["B"], c["H"], c["W"])
        kern_ms = {
            "volume_gwc_concat": time_kernel(torch, lambda: ops.combined_volume(dev["fl"], dev["fr"], dev["pl"], dev["pr"], D, G), 20, flush),
            "corr_pyramid": time_kernel(torch, lambda: ops.corr_pyramid(dev["fl"], dev["fr"], L, FSB_PREC_TF32X3), 20, flush),
            "corr_pyramid_fp32_cuda_cores": time_kernel(torch, lambda: ops.corr_pyramid(dev["fl"], dev["fr"], L, FSB_PREC_FP32), 20, flush),
            "geo_pyramid": time_kernel(torch, lambda: ops.geo_pyramid(dev["volume"], L), 20, flush),
            "geo_lookup": time_kernel(torch, lambda: ops.geo_lookup_nhwc16(geo.geo_levels, geo.corr_levels, dev["disp"], coords, r, plan_b.corr16), 20, flush),
            "geo_lookup_fp32_nchw": time_kernel(torch, lambda: geo(dev["disp"], coords), 20, flush),
        }
        # the steps either side of the loop: not part of the timed step, reported beside it
        cls_logits = torch.randn(c["B"], D, c["H"], c["W"], device=device)
        spx_logits = torch.randn(c["B"], 9, 4 * c["H"], 4 * c["W"], device=device)
        kern_ms["softmax_disparity_regression"] = time_kernel(torch, lambda: ops.disparity_regression(cls_logits, True), 20, flush)
        kern_ms["softmax_context_upsample"] = time_kernel(torch, lambda: ops.context_upsample(dev["disp"], spx_logits, True, 4.0), 20, flush)
        per_step = {"volume_gwc_concat": 1, "corr_pyramid": 1, "corr_pyramid_fp32_cuda_cores": 0, "geo_pyramid": 1,
                    "geo_lookup": c["iters"], "geo_lookup_fp32_nchw": 0, "softmax_disparity_regression": 0,
                    "softmax_context_upsample": 0}
        alg = algorithmic_bytes(c)
        px_ = c["B"] * c["H"] * c["W"]
        alg["softmax_disparity_regression"] = px_ * (D + 1) * 4
        alg["softmax_context_upsample"] = px_ * (4 + 16 * (9 + 1) * 4)
        alg["corr_pyramid_fp32_cuda_cores"] = alg["corr_pyramid"]
        alg["geo_lookup_fp32_nchw"] = alg["geo_lookup"]
        # the fused lookup writes 16-bit NHWC rows of Cpad channels instead of fp32 NCHW
        K_ = 2 * r + 1
        cpad = (L * K_ * (c["Cg"] + 1) + 63) // 64 * 64
        alg["geo_lookup"] = c["B"] * c["H"] * c["W"] * (4 + L * (c["Cg"] + 1) * (K_ + 1) * 4 + cpad * 2)
        kernels = {}
        for k, ms in kern_ms.items():
            gbs = alg[k] / (ms * 1e-3) / 1e9
            kernels[k] = {"ms": ms, "launches_per_step": per_step[k], "algorithmic_bytes": alg[k],
                          "achieved_gbs": gbs, "frac_of_measured_hbm": gbs / peaks["hbm_gbs"],
                          "share_of_step": ms * per_step[k] / (ms_total / args.steps)}
        # the gwc half of kernel 1 is fp32-FMA bound, not HBM bound: C/G MACs per output element
        gwc_macs = c["B"] * c["H"] * c["W"] * c["D"] * c["G"] * (c["C"] // c["G"])
        kernels["volume_gwc_concat"]["fp32_fma_bound_ms"] = gwc_macs / (148 * 128 * 1.965e9) * 1e3
        top = max(kernels, key=lambda k: kernels[k]["share_of_step"])
        roofline = {"kernel": top, "bound": "hbm", "achieved": kernels[top]["achieved_gbs"], "peak": peaks["hbm_gbs"],
                    "unit": "GB/s", "frac": kernels[top]["frac_of_measured_hbm"], "traffic": None,
                    "peak_source": peaks["source"] + " (MEASURED_PEAKS.json burst copy)", "kernels": kernels}
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(traffic_file):  # DRAM bytes per launch from the committed ncu --set full captures
            with open(traffic_file) as fh:
                tj = json.load(fh)
            roofline["traffic"] = tj.get(top)
            for k in kernels:
                kernels[k]["traffic"] = tj.get(k)

        # ---- CPU baseline (oracle port) on the host cores, bounded sample --------------------
        threads = os.cpu_count() or 1
        if world == 1:
            best = min(cpu_sample(60, 2, threads)[0] for _ in range(2))
            cpu = {"value": 1.0 / best, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "rows 0..59 of 120 (full width, D=104): volumes + geo init + 2 of 32 iterations, best of 2, "
                             "extrapolated linearly to one pair; torch fp32 CPU oracle (oracle/hotpath.py)"}
        else:
            cpu = {"value": None, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "timed at N=1 only (python bench.py --gpus 1)"}

        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f16 operands / f32 accumulate (update block); f32 (volumes, geo)",
            "data": "synthetic",
            "config": {"workload": workload_name(), "pairs_per_gpu_per_step": c["B"], "iters": c["iters"],
                       "l2": "inputs+outputs of a step (224 MB filtered volume, 256 MB cost volume, 335 MB geo pyramid) "
                             "exceed the 126 MB L2; per-kernel timings flush L2 with a 512 MiB write before each launch",
                       "parallelism": f"pairs sharded over {world} GPU(s), no collective on the path",
                       "update_block": "tcgen05 implicit-GEMM convolutions with fused gate epilogues (csrc/conv.cu), one CUDA graph per step"},
            "us_per_gru_iter": None,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches),
            "roofline": roofline,
            "cpu_baseline": cpu,
            "clocks": clocks.summary(tw0, te1),
            "wall_ms_per_step": 1e3 * (tw1 - tw0) / args.steps,
        }
        # us per GRU iteration: the refinement loop alone (one graph replay per iteration)
        ws = fsb.Combined_Geo_Encoding_Volume.make_workspace(c["B"], c["Cg"], D, c["H"], c["W"], c["W"], L, device)
        geo_ws = fsb.Combined_Geo_Encoding_Volume(dev["fl"], dev["fr"], dev["volume"], num_levels=L, dx=hp.dx, workspace=ws)
        ub = hp.update_block
        for _ in range(2):
            ub.refine(geo_ws, dev["net"], dev["inp"], dev["att"], dev["disp"], 2)
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        g0.record()
        ub.refine(geo_ws, dev["net"], dev["inp"], dev["att"], dev["disp"], c["iters"])
        g1.record()
        torch.cuda.synchronize()
        us_iter = 1e3 * g0.elapsed_time(g1) / c["iters"]
        line["us_per_gru_iter"] = us_iter

        # ---- tensor roofline of the dominant kernel (conv_umma_kernel: every convolution / linear of the
        # update block).  Per-launch CUDA-event durations of one step issued launch by launch on one
        # stream; achieved = algorithmic FLOPs (2*px*Cout*Cin*k*k, unpadded) / sum of durations.
        plan = ub.plan_for(c["B"], c["H"], c["W"])
        steps = [(n_, f_) for n_, f_ in zip(plan.names, plan.steps) if callable(f_)]
        reps = 10
        tot = [0.0] * len(steps)
        for _ in range(reps):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(steps) + 1)]
            ev[0].record()
            for i, (_, fn) in enumerate(steps):
                fn()
                ev[i + 1].record()
            torch.cuda.synchronize()
            for i in range(len(steps)):
                tot[i] += ev[i].elapsed_time(ev[i + 1]) / reps
        conv_ms = sum(t for (n_, f_), t in zip(steps, tot) if hasattr(f_, "flops"))
        other_ms = sum(t for (n_, f_), t in zip(steps, tot) if not hasattr(f_, "flops"))
        conv_flops = synthetic.update_block_flops(state, c["B"], c["H"], c["W"])
        n_conv = sum(1 for n_, f_ in steps if hasattr(f_, "flops"))
        tf = conv_flops / (conv_ms * 1e-3) / 1e12
        peak_tf = peaks["bf16_tflops_sustained"]
        conv_share = (conv_ms * c["iters"]) / (ms_total / args.steps)
        hbm = line["roofline"]
        line["roofline"] = {
            "kernel": "conv_umma_kernel (tcgen05 implicit-GEMM convolution, all %d launches of one update step)" % n_conv,
            "bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf,
            "traffic": None,
            "peak_source": peaks["source"] + " (MEASURED_PEAKS.json bf16_tflops_sustained: the kernel runs inside a long step)",
            "algorithmic_flops_per_step": conv_flops, "launches_per_step": n_conv * c["iters"],
            "ms_per_update_step_sum_of_launches": conv_ms, "ms_other_kernels_per_update_step": other_ms,
            "share_of_step": min(conv_share, 1.0),
            "whole_iteration_tflops": conv_flops / (us_iter * 1e-6) / 1e12,
            "hbm_kernels": hbm["kernels"],
        }
        if os.path.exists(traffic_file):
            with open(traffic_file) as fh:
                line["roofline"]["traffic"] = json.load(fh).get("conv_umma_kernel")
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_REAL_STDOUT = None


def _quiet_stdout():
    """Route fd 1 to stderr for the rest of the run: libraries (NCCL prints its version line on stdout) must not
    get in front of the ONE JSON line, which emit() writes to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(obj) -> None:
    data = (json.dumps(obj) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--pairs", type=int, default=8, help="stereo pairs per GPU per step (one batch through the hot path)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    args = ap.parse_args()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 or args.gpus == 1:
        _quiet_stdout()  # (the --gpus N re-launcher keeps its stdout: its children write the line)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
