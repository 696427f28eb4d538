#!/usr/bin/env python
"""bench.py -- FoundationStereo hot path on B200: pairs/s @ 480x640, max_disp 416, 32 GRU iterations.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config NAME] [--pairs P]

A "step" is one pass of the hot path over one batch of synthetic stereo pairs per GPU:
    volumes (kernel 1) -> geometry-encoding volume (kernels 2, 2b) -> 32 x (kernel 3 + update block).
The backbone and the 3-D hourglass of the reference are out of scope (BASELINE.json), so the
tensors they would produce (features, filtered volume, context) are seeded synthetic inputs.

--config (BASELINE.json configs; the default is the one the metric is quoted on and the driver runs):
    vits480     configs[1]  ViT-S 480x640 -> 1/4-res 120x160, C=128, D=104; weak scaling, --pairs per GPU per step
    vitl_hiera  configs[2]  ViT-L 1024x1536 --hiera 1: pass A 128x192 then pass B 256x384 (C=224), the coarse
                            result initialises the fine pass (core/foundation_stereo.py:257-274)
    kitti64     configs[3]  64 KITTI-shape pairs (96x312) sharded over the GPUs: STRONG scaling, micro-batches of 8
    erp         configs[4]  ERP panorama 1024x2048 -> 256x512 (wide-W init correlation)

One JSON line on stdout (rank 0).  `value` = pairs/s with inputs resident in HBM;
`e2e` = the same through the public API with pinned host buffers, H2D + D2H inside the
timed region; `roofline` = the dominant kernel of libfsb200.so against MEASURED_PEAKS.json;
`cpu_baseline` = the reference's OWN PyTorch code for the path (oracle/_ref, carried to the box by
tools/make_ref.py) on the host cores, one whole pair, not extrapolated.
`--impl reference` times that CPU path alone (rank 0 only): every step is one whole pair.
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

METRIC = "pairs/sec @480x640 D416 32it"
UNIT = "pairs/s"
COMMON = dict(D=104, Cp=12, Cg=28, G=8, L=2, r=4, iters=32)
CONFIGS = {
    "vits480": dict(COMMON, passes=[dict(C=128, H=120, W=160)], pairs=8, scaling="weak",
                    workload="hot path of FoundationStereo ViT-S config (BASELINE configs[1]): 1/4-res 120x160, C=128, "
                             "max_disp=416 (D=104), corr_levels=2, radius=4, 32 x (lookup + update block); a step = one "
                             "batch of pairs per GPU (config.pairs_per_gpu_per_step)"),
    "vitl_hiera": dict(COMMON, passes=[dict(C=224, H=128, W=192), dict(C=224, H=256, W=384)], pairs=2, scaling="weak",
                       workload="hot path of the ViT-L hierarchical run (BASELINE configs[2], 1024x1536 --hiera 1): pass A "
                                "at 1/4-res 128x192 then pass B at 256x384 (C=224, D=104, 32 iterations each), pass B "
                                "initialised from pass A's disparity (core/foundation_stereo.py:257-274)"),
    "kitti64": dict(COMMON, passes=[dict(C=128, H=96, W=312)], pairs=8, total_pairs=64, scaling="strong",
                    workload="64 KITTI-shape pairs (BASELINE configs[3], 375x1242 -> 1/4-res 96x312, C=128, D=104, 32 "
                             "iterations) sharded over the GPUs in contiguous ranges (shard.run_sharded), micro-batches "
                             "of 8 pairs; a step = all 64 pairs"),
    "erp": dict(COMMON, passes=[dict(C=128, H=256, W=512)], pairs=2, scaling="weak",
                workload="ERP panorama (BASELINE configs[4], 1024x2048 -> 1/4-res 256x512, C=128, D=104, 32 iterations): "
                         "the wide-W init correlation (268 MB per pair)"),
}


def config_block(name, pairs, world):
    """The `config` object of the JSON line -- identical for the b200 and the reference arm."""
    c = CONFIGS[name]
    return {"workload": c["workload"], "name": name, "pairs_per_gpu_per_step": pairs, "iters": c["iters"],
            "l2": "inputs+outputs of a step (per pair: 112 MB fp16 filtered volume, 256 MB cost volume, 335 MB geo pyramid) "
                  "exceed the 126 MB L2; per-kernel timings flush L2 with a 512 MiB write before each launch",
            "parallelism": "pairs sharded over the GPUs, no collective on the path",
            "update_block": "tcgen05 implicit-GEMM convolutions with fused gate epilogues (csrc/conv.cu), ONE CUDA graph for the whole refinement loop"}


# --------------------------------------------------------------------------------------
# CPU arm: the reference's own code for the path on the host cores -- one WHOLE pair per step
# --------------------------------------------------------------------------------------
class CpuReference:
    """`oracle/_ref` (the unmodified reference, kind "reference") when it travelled to this box, else the
    oracle port (kind "port").  One call = one whole pair of the config's LAST pass shape... every pass for
    the hierarchical config: volumes + geo init + 32 x (lookup + update block), fp32, all host threads."""

    def __init__(self, name: str):
        import torch
        from foundationstereo_b200 import synthetic
        self.torch = torch
        self.c = CONFIGS[name]
        self.threads = os.cpu_count() or 1
        torch.set_num_threads(self.threads)
        c = self.c
        self.state = synthetic.random_update_block_state(seed=0, corr_levels=c["L"], corr_radius=c["r"],
                                                         volume_dim=c["Cg"])
        self.x = [synthetic.hot_path_inputs(1, p["C"], p["H"], p["W"], c["D"], c["Cp"], c["Cg"], seed=0)
                  for p in c["passes"]]
        for x in self.x:   # fp16-representable values, as the model's autocast region produces them
            x["volume"] = x["volume"].half().float()
        from oracle import refmodel
        if refmodel.available():
            self.kind = "reference"
            self.rm = refmodel
            self.ub = refmodel.make_reference_update_block(self.state, c["L"], c["r"])
            self.what = ("the reference's own build_gwc_volume / build_concat_volume (core/submodule.py:399-427), "
                         "Combined_Geo_Encoding_Volume (core/geometry.py:18-65) and BasicSelectiveMultiUpdateBlock "
                         "(core/update.py:141-159), unmodified, from oracle/_ref")
        else:
            self.kind = "port"
            import oracle
            self.oracle = oracle
            self.what = "oracle/hotpath.py (CPU restatement of the reference; oracle/_ref is absent on this box)"

    def pair(self):
        """-> (seconds, per-stage seconds) for one pair through every pass of the config."""
        c, torch = self.c, self.torch
        t0 = time.perf_counter()
        stages = {}
        for x in self.x:
            if self.kind == "reference":
                _, t = self.rm.reference_hot_path_cpu(self.ub, x, 4 * c["D"], c["G"], c["L"], c["r"], c["iters"])
            else:
                o = self.oracle
                with torch.no_grad():
                    a = time.perf_counter()
                    o.gwc_volume(x["fl"], x["fr"], c["D"], c["G"])
                    o.concat_volume(x["pl"], x["pr"], c["D"])
                    b = time.perf_counter()
                    geo = o.GeoEncodingVolume(x["fl"], x["fr"], x["volume"], c["L"], c["r"])
                    d = time.perf_counter()
                    o.refine(self.state, geo, x["net"], x["inp"], x["att"], x["disp"], c["iters"])
                    e = time.perf_counter()
                t = dict(volumes_s=b - a, geo_init_s=d - b, loop_s=e - d, total_s=e - a)
            for k, v in t.items():
                stages[k] = stages.get(k, 0.0) + v
        return time.perf_counter() - t0, stages

    def sample_text(self):
        shapes = " + ".join(f"{p['H']}x{p['W']} C={p['C']}" for p in self.c["passes"])
        return (f"ONE whole pair per step, not extrapolated: 1/4-res {shapes}, D={self.c['D']}, volumes + geo init + "
                f"{self.c['iters']} x (lookup + update block); {self.what}; torch fp32, {self.threads} threads")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cpu = CpuReference(args.config)
    for _ in range(args.warmup):
        cpu.pair()
    times, stages = [], []
    for _ in range(args.steps):
        t, st = cpu.pair()
        times.append(t)
        stages.append(st)
    mean_s = statistics.mean(times)
    value = 1.0 / mean_s
    pairs = args.pairs or CONFIGS[args.config]["pairs"]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * mean_s, "higher_is_better": True,
        "scaling": CONFIGS[args.config]["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_block(args.config, pairs, int(os.environ.get("WORLD_SIZE", "1"))),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cpu.threads, "kind": cpu.kind, "sample": cpu.sample_text(),
                         "pair_seconds_min_max": [min(times), max(times)],
                         "stage_seconds_mean": {k: statistics.mean(s[k] for s in stages) for k in stages[0]}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU arm: a step here is ONE pair (the GPU arm's step is config.pairs_per_gpu_per_step pairs); "
                "the metric, pairs/s, is per pair in both",
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
    cpad = (L * K * (Cg + 1) + 63) // 64 * 64
    return {
        "volume_gwc_concat": px * ((2 * C + 2 * Cp) * 4 + (G + 2 * Cp) * D * 4),
        "corr_pyramid": 8 * px * C + 4 * px * sum(W >> l for l in range(L)),
        "geo_pyramid": px * Cg * (c.get("vol_bytes", 4) * D + 4 * sum(D >> l for l in range(L))),
        # the fused lookup writes 16-bit NHWC rows of Cpad channels instead of fp32 NCHW
        "geo_lookup": px * (4 + L * (Cg + 1) * (K + 1) * 4 + cpad * 2),
        "geo_lookup_fp32_nchw": px * (4 + L * (Cg + 1) * (K + 1) * 4 + L * K * (Cg + 1) * 4),
        "softmax_disparity_regression": px * (D + 1) * 4,
        "softmax_context_upsample": px * (4 + 16 * (9 + 1) * 4),
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
               "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup", str(args.warmup),
               "--config", args.config] + (["--pairs", str(args.pairs)] if args.pairs else [])
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
    from foundationstereo_b200 import ops, shard, synthetic
    from foundationstereo_b200._lib import FSB_PREC_FP32, FSB_PREC_TF32X3

    cfg = CONFIGS[args.config]
    pairs = args.pairs or cfg["pairs"]      # pairs per launch of the path on one GPU (the micro-batch)
    iters = cfg["iters"]
    strong = cfg["scaling"] == "strong"
    if strong:
        lo, hi = shard.shard_range(cfg["total_pairs"], rank, world)
        micro = list(shard.micro_batches(lo, hi, pairs))
        local_pairs_per_step = hi - lo
    else:
        micro = [(0, pairs)]
        local_pairs_per_step = pairs
    passes = [dict(COMMON, **p, B=pairs) for p in cfg["passes"]]
    state = synthetic.random_update_block_state(seed=0, corr_levels=cfg["L"], corr_radius=cfg["r"], volume_dim=cfg["Cg"])
    hp = fsb.StereoHotPath(state, max_disp=4 * cfg["D"], corr_levels=cfg["L"], corr_radius=cfg["r"], num_groups=cfg["G"],
                           device=device)
    hp.record_loop = True

    def pin(t):
        return t.contiguous().pin_memory()
    # one micro-batch of `pairs` distinct synthetic pairs per pass (a strong-scaling step re-uses it for every
    # micro-batch of the rank's range: the timing does not depend on the values)
    xs = [synthetic.hot_path_inputs(pairs, c["C"], c["H"], c["W"], c["D"], c["Cp"], c["Cg"], seed=rank + 17 * i)
          for i, c in enumerate(passes)]
    # the filtered cost volume leaves the reference's autocast region in fp16 and only becomes fp32 through the caller's
    # `.float()` (core/foundation_stereo.py:213-229): it is handed over (and uploaded) as fp16 and kernel 2b reads it as
    # such (fsb_geo_pyramid_in); the CPU arm gets the same values through `.float()`
    for x in xs:
        x["volume"] = x["volume"].half()
    hosts = [{k: ([pin(t) for t in v] if isinstance(v, list) else pin(v)) for k, v in x.items()} for x in xs]
    del xs
    devs = [{k: ([t.to(device, non_blocking=True) for t in v] if isinstance(v, list) else v.to(device, non_blocking=True))
             for k, v in h.items()} for h in hosts]
    h2d_per_micro = sum(sum(t.numel() * t.element_size() for t in v) if isinstance(v, list)
                        else v.numel() * v.element_size() for h in hosts for v in h.values())

    def run_passes(ds):
        """all passes of the config over one micro-batch: a later pass starts from the earlier pass's result,
        up-scaled like run_hierachical does (x2 in size, x2 in value; core/foundation_stereo.py:263-271)."""
        disp = None
        for i, d in enumerate(ds):
            init = d["disp"]
            if disp is not None:
                init = torch.nn.functional.interpolate(disp, size=init.shape[-2:], mode="bilinear", align_corners=True) * 2.0
            disp, _, _ = hp.run(d["fl"], d["fr"], d["pl"], d["pr"], d["volume"], init, d["net"], d["inp"], d["att"], iters)
        return disp

    def step():
        out = None
        for _ in micro:
            out = run_passes(devs)
        return out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident-input timing ---------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        out = step()
    barrier()
    clocks = Clocks(local_rank)
    clocks.start()
    time.sleep(0.3)
    hp.loop_events.clear()
    n0 = fsb.launch_count() + hp.update_block.replayed_kernels
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    tw0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        out = step()
    e1.record()
    barrier()
    tw1 = time.perf_counter()
    launches = fsb.launch_count() + hp.update_block.replayed_kernels - n0  # eager launches + kernels inside graph replays
    ms_total = e0.elapsed_time(e1)
    ms_loops = sum(a.elapsed_time(b) for a, b in hp.loop_events)   # the 32-iteration loops inside the timed steps
    n_loops = len(hp.loop_events)
    hp.record_loop = False

    # ---- end to end: pinned host buffers in, disparity out, every step -------------------
    # single-pass configs: StereoHotPath.run_pinned (the upload of batch i+1 runs on a copy stream while batch i
    # computes); the hierarchical config uploads and runs its passes in order.  Every micro-batch's inputs cross
    # PCIe inside the timed region and every micro-batch's disparity comes back.
    n_micro = len(micro)
    out_hosts = [torch.empty_like(out, device="cpu").pin_memory() for _ in range(max(args.steps * n_micro, 2))]

    def e2e(nsteps):
        if len(hosts) == 1:
            hp.run_pinned([hosts[0]] * (nsteps * n_micro), iters, out_hosts[:nsteps * n_micro])
            return
        for s in range(nsteps):
            ds = [{k: ([t.to(device, non_blocking=True) for t in v] if isinstance(v, list) else v.to(device, non_blocking=True))
                   for k, v in h.items()} for h in hosts]
            out_hosts[s].copy_(run_passes(ds), non_blocking=True)
        torch.cuda.synchronize()

    e2e(1)
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    te0 = time.perf_counter()
    f0.record()
    e2e(args.steps)
    f1.record()
    barrier()
    te1 = time.perf_counter()
    ms_e2e = max(f0.elapsed_time(f1), 1e3 * (te1 - te0))
    d2h_bytes = out_hosts[0].numel() * out_hosts[0].element_size() * n_micro

    # whole-job figures: SUM of pairs, MAX of device times over ranks (foundationstereo_b200/shard.py)
    total_pairs, ms_total, value = shard.reduce_throughput(local_pairs_per_step * args.steps, ms_total, device)
    _, ms_e2e, e2e_value = shard.reduce_throughput(local_pairs_per_step * args.steps, ms_e2e, device)

    if rank == 0:
        peaks = measured_peaks()
        c = passes[-1]                      # per-kernel figures: the (largest) last pass
        dev = devs[-1]
        D, G, L, r = c["D"], c["G"], c["L"], c["r"]
        ub = hp.update_block
        ms_step = ms_total / args.steps

        # ---- inside the clocks window: the refinement loop alone and the per-launch durations of one update step
        ws = fsb.Combined_Geo_Encoding_Volume.make_workspace(c["B"], c["Cg"], D, c["H"], c["W"], c["W"], L, device)
        geo_ws = fsb.Combined_Geo_Encoding_Volume(dev["fl"], dev["fr"], dev["volume"], num_levels=L, dx=hp.dx, workspace=ws)
        for _ in range(2):   # eager pass, then the capture of the whole-loop graph for this geometry volume
            ub.refine(geo_ws, dev["net"], dev["inp"], dev["att"], dev["disp"], iters)
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        g0.record()
        ub.refine(geo_ws, dev["net"], dev["inp"], dev["att"], dev["disp"], iters)
        g1.record()
        torch.cuda.synchronize()
        us_iter_alone = 1e3 * g0.elapsed_time(g1) / iters
        plan = ub.plan_for(c["B"], c["H"], c["W"])
        steps_ = [(n_, f_) for n_, f_ in zip(plan.names, plan.steps) if callable(f_)]
        # time encoder.convc1 as the captured loop runs it (with the lookup fused in, if enabled)
        plan.loop_active = bool(ub.fused_lookup_conv and ub.pipelined and plan.fused_c1 is not None)
        reps = 10
        tot = [0.0] * len(steps_)
        for _ in range(reps):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(steps_) + 1)]
            ev[0].record()
            for i, (_, fn) in enumerate(steps_):
                fn()
                ev[i + 1].record()
            torch.cuda.synchronize()
            for i in range(len(steps_)):
                tot[i] += ev[i].elapsed_time(ev[i + 1]) / reps
        plan.loop_active = False
        t_gpu_end = time.perf_counter()
        clocks.stop()
        # with the lookup fused into encoder.convc1 (csrc/lookconv.cu, the default loop) that launch is the HBM-bound gather
        # with a small GEMM under it: it is reported under hbm_kernels, not in the tensor-core roofline
        fused_c1 = bool(ub.fused_lookup_conv and ub.pipelined and plan.fused_c1 is not None)
        is_conv = lambda n_, f_: hasattr(f_, "flops") and not (fused_c1 and n_ == "encoder.convc1")
        conv_ms = sum(t for (n_, f_), t in zip(steps_, tot) if is_conv(n_, f_))
        other_ms = sum(t for (n_, f_), t in zip(steps_, tot) if not is_conv(n_, f_))
        n_conv = sum(1 for n_, f_ in steps_ if is_conv(n_, f_))
        flops_by_pass = [synthetic.update_block_flops(state, p["B"], p["H"], p["W"]) for p in passes]
        c1_flops = sum(f_.flops for n_, f_ in steps_ if n_ == "encoder.convc1") if fused_c1 else 0.0
        c1_ms = sum(t for (n_, f_), t in zip(steps_, tot) if n_ == "encoder.convc1")
        conv_flops = flops_by_pass[-1] - c1_flops
        tf = conv_flops / (conv_ms * 1e-3) / 1e12
        peak_tf = peaks["bf16_tflops_sustained"]
        # in-step figure: the loops were timed with CUDA events INSIDE the timed steps (StereoHotPath.record_loop)
        loop_flops_per_step = iters * sum(flops_by_pass) * n_micro
        ms_loops_per_step = ms_loops / args.steps
        in_step_tf = loop_flops_per_step / (ms_loops_per_step * 1e-3) / 1e12
        whole_step_tf = loop_flops_per_step / (ms_step * 1e-3) / 1e12
        us_iter = 1e3 * ms_loops / (n_loops * iters) if len(passes) == 1 else None

        # ---- per-kernel device times for the HBM-bound kernels (L2 flushed before each launch) ----------------
        flush = torch.empty(512 << 20, dtype=torch.uint8, device=device)
        geo = fsb.Combined_Geo_Encoding_Volume(dev["fl"], dev["fr"], dev["volume"], num_levels=L, dx=hp.dx)
        coords = torch.arange(c["W"], dtype=torch.float32, device=device).reshape(1, 1, -1, 1).repeat(c["B"], c["H"], 1, 1)
        kern_ms = {
            "volume_gwc_concat": time_kernel(torch, lambda: ops.combined_volume(dev["fl"], dev["fr"], dev["pl"], dev["pr"], D, G), 20, flush),
            "corr_pyramid": time_kernel(torch, lambda: ops.corr_pyramid(dev["fl"], dev["fr"], L, FSB_PREC_TF32X3), 20, flush),
            "geo_pyramid": time_kernel(torch, lambda: ops.geo_pyramid(dev["volume"], L), 20, flush),
            "geo_lookup": time_kernel(torch, lambda: ops.geo_lookup_nhwc16(geo.geo_levels, geo.corr_levels, dev["disp"], coords, r, plan.corr16), 20, flush),
            "geo_lookup_fp32_nchw": time_kernel(torch, lambda: geo(dev["disp"], coords), 20, flush),
        }
        per_step = {"volume_gwc_concat": n_micro, "corr_pyramid": n_micro, "geo_pyramid": n_micro,
                    "geo_lookup": 0 if fused_c1 else iters * n_micro, "geo_lookup_fp32_nchw": 0}
        if fused_c1:   # timed above among the loop's launches (no L2 flush: the 2.7 GB geometry pyramid does not fit anyway)
            kern_ms["lookup_convc1"] = c1_ms
            per_step["lookup_convc1"] = iters * n_micro
        # opt-in banded init correlation (disparities in [0, D]): reported beside the full-matrix default
        bw0 = ops.corr_band_geometry(D, r, L)[0]
        if bw0 + (4 << (L - 1)) <= 224:
            band_lv = [torch.empty((c["B"] * c["H"] * c["W"], bw0 >> l), dtype=torch.float32, device=device) for l in range(L)]
            kern_ms["corr_pyramid_band"] = time_kernel(
                torch, lambda: ops.corr_pyramid_band(dev["fl"], dev["fr"], L, FSB_PREC_TF32X3, D, r, band_lv), 20, flush)
            kern_ms["geo_lookup_band"] = time_kernel(
                torch, lambda: ops.geo_lookup_band(geo.geo_levels, band_lv, dev["disp"], coords, r, D, plan.corr16), 20, flush)
            per_step.update(corr_pyramid_band=0, geo_lookup_band=0)
        if args.config == "vits480":
            kern_ms["corr_pyramid_fp32_cuda_cores"] = time_kernel(torch, lambda: ops.corr_pyramid(dev["fl"], dev["fr"], L, FSB_PREC_FP32), 20, flush)
            # the steps either side of the loop: not part of the timed step, reported beside it
            cls_logits = torch.randn(c["B"], D, c["H"], c["W"], device=device)
            spx_logits = torch.randn(c["B"], 9, 4 * c["H"], 4 * c["W"], device=device)
            kern_ms["softmax_disparity_regression"] = time_kernel(torch, lambda: ops.disparity_regression(cls_logits, True), 20, flush)
            kern_ms["softmax_context_upsample"] = time_kernel(torch, lambda: ops.context_upsample(dev["disp"], spx_logits, True, 4.0), 20, flush)
            per_step.update(corr_pyramid_fp32_cuda_cores=0, softmax_disparity_regression=0, softmax_context_upsample=0)
        alg = algorithmic_bytes(dict(c, vol_bytes=dev["volume"].element_size()))
        alg["corr_pyramid_fp32_cuda_cores"] = alg["corr_pyramid"]
        px_ = c["B"] * c["H"] * c["W"]
        alg["corr_pyramid_band"] = 8 * px_ * c["C"] + 4 * px_ * sum(bw0 >> l for l in range(L))
        K_ = 2 * r + 1
        alg["geo_lookup_band"] = alg["geo_lookup"]   # the lookup reads the same windows, only their addresses differ
        # lookup fused into convc1: disparity + coordinate, the (K+1)-row windows of every level, convc1's 256 x 16-bit output
        alg["lookup_convc1"] = px_ * (8 + L * (c["Cg"] + 1) * (K_ + 1) * 4 + 256 * 2)
        kernels = {}
        for k, ms in kern_ms.items():
            gbs = alg[k] / (ms * 1e-3) / 1e9
            kernels[k] = {"ms": ms, "launches_per_step": per_step[k], "algorithmic_bytes": alg[k],
                          "achieved_gbs": gbs, "frac_of_measured_hbm": gbs / peaks["hbm_gbs"],
                          "share_of_step": ms * per_step[k] / ms_step}
        # the gwc half of kernel 1 is fp32-FMA bound, not HBM bound: C/G MACs per output element
        gwc_macs = c["B"] * c["H"] * c["W"] * c["D"] * c["G"] * (c["C"] // c["G"])
        kernels["volume_gwc_concat"]["fp32_fma_bound_ms"] = gwc_macs / (148 * 128 * 1.965e9) * 1e3
        kernels["corr_pyramid"]["tf32x3_flops"] = 3 * 2.0 * c["B"] * c["H"] * c["W"] * c["W"] * c["C"]
        traffic = {}
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(traffic_file) and args.config == "vits480" and pairs == 8:
            with open(traffic_file) as fh:  # DRAM bytes per launch from the committed ncu --set full captures (8 pairs)
                traffic = json.load(fh)
            for k in kernels:
                kernels[k]["traffic"] = traffic.get(k)

        # ---- CPU baseline: the reference's own code, one whole pair, N=1 only ------------------------------
        threads = os.cpu_count() or 1
        if world == 1 and not args.no_cpu:
            cpu_ref = CpuReference(args.config)
            t_pair, stages = cpu_ref.pair()
            cpu = {"value": 1.0 / t_pair, "unit": UNIT, "cores": cpu_ref.threads, "kind": cpu_ref.kind,
                   "sample": cpu_ref.sample_text() + "; timed once (no warm-up)", "pair_seconds": t_pair,
                   "stage_seconds": stages}
        else:
            cpu = {"value": None, "unit": UNIT, "cores": threads, "kind": "reference",
                   "sample": "timed at N=1 only (python bench.py --gpus 1)"}

        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": cfg["scaling"], "vs_baseline": None,
            "dtype": "f16 operands / f32 accumulate (update block); f32 (volumes, geo)", "data": "synthetic",
            "config": config_block(args.config, pairs, world),
            "pairs_per_step_all_gpus": total_pairs // args.steps,
            "us_per_gru_iter": us_iter, "us_per_gru_iter_loop_alone": us_iter_alone,
            "ms_outside_loop": ms_step - ms_loops_per_step,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_per_micro * n_micro,
                    "d2h_bytes_per_step": d2h_bytes, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches),
            "roofline": {
                "kernel": "conv_umma_kernel (tcgen05 implicit-GEMM convolution, all %d tensor-core launches of one update step)" % n_conv,
                "bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf,
                "traffic": traffic.get("conv_umma_kernel"),
                "peak_source": peaks["source"] + " (MEASURED_PEAKS.json bf16_tflops_sustained: the kernel runs inside a long step)",
                "algorithmic_flops_per_update_step": conv_flops, "launches_per_step": n_conv * iters * n_micro,
                "ms_per_update_step_sum_of_launches": conv_ms, "ms_other_kernels_per_update_step": other_ms,
                # the same FLOPs over the loop time measured INSIDE the timed steps (all 29 kernels of an iteration, two
                # lanes): what the tensor launches deliver in the step, the figure `frac` must agree with
                "in_step_achieved": in_step_tf, "in_step_frac": in_step_tf / peak_tf,
                "whole_step_achieved": whole_step_tf, "whole_step_frac": whole_step_tf / peak_tf,
                "share_of_step": min(ms_loops_per_step / ms_step, 1.0),
                "per_launch_ms": {n_: t for (n_, f_), t in zip(steps_, tot)},
                "hbm_kernels": kernels,
            },
            "cpu_baseline": cpu,
            "clocks": clocks.summary(tw0, t_gpu_end),
            "wall_ms_per_step": 1e3 * (tw1 - tw0) / args.steps,
        }
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
    if os.environ.get("FSB_BENCH_WATCHDOG"):  # debugging aid: dump every thread's Python stack and exit after N seconds
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["FSB_BENCH_WATCHDOG"]), exit=True)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--pairs", type=int, default=0, help="stereo pairs per launch of the path on one GPU (default: per config)")
    ap.add_argument("--config", default="vits480", choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg of the b200 arm (profiling runs)")
    args = ap.parse_args()
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 or args.gpus == 1:
        _quiet_stdout()  # (the --gpus N re-launcher keeps its stdout: its children write the line)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
