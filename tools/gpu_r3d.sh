#!/bin/bash
# r3d: two GPUs of one box, launched as the driver does: weak scaling at HEAD, KITTI strong scaling, the reference arm under torchrun
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 8 --warmup 3 > gpurun_out/r3d_bench_2gpu.json 2> gpurun_out/r3d_bench_2gpu.err; echo "2gpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --config kitti64 --steps 3 --warmup 3 > gpurun_out/r3d_bench_kitti64_2gpu.json 2> gpurun_out/r3d_bench_kitti64_2gpu.err; echo "kitti 2gpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/r3d_bench_ref_2gpu.json 2> gpurun_out/r3d_bench_ref_2gpu.err; echo "ref 2gpu rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3d_*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d.get("impl", "b200"), round(d["value"], 3), d["unit"], "n_gpus", d["n_gpus"], "e2e", d.get("e2e", {}).get("value"), d.get("clocks"))
    except Exception as e:
        print(f, "unreadable", e)
PY
