#!/bin/bash
# 2-GPU pass: default weak-scaling line and the KITTI x 64 strong-scaling line through shard.shard_range / micro_batches
mkdir -p gpurun_out
nvidia-smi -L
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "16bit_volume or golden" 2>&1 | tail -1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/r2m_bench_2gpu.json 2> gpurun_out/r2m_bench_2gpu.err; echo "weak 2gpu rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 --config kitti64 > gpurun_out/r2m_bench_kitti64_2gpu.json 2> gpurun_out/r2m_bench_kitti64_2gpu.err; echo "kitti64 2gpu rc=$?"
timeout 600 python bench.py --steps 3 --warmup 3 --config kitti64 --no-cpu > gpurun_out/r2m_bench_kitti64_1gpu.json 2> gpurun_out/r2m_bench_kitti64_1gpu.err; echo "kitti64 1gpu rc=$?"
for f in r2m_bench_2gpu r2m_bench_kitti64_2gpu r2m_bench_kitti64_1gpu; do python - <<PY
import json
try:
    b=json.load(open("gpurun_out/$f.json")); print("$f", b["n_gpus"], round(b["value"],2), "e2e", round(b["e2e"]["value"],2), "ms/step", round(b["ms_per_step"],1), b["scaling"], b.get("pairs_per_step_all_gpus"))
except Exception as e: print("$f ERR", e)
PY
done
tail -3 gpurun_out/r2m_bench_2gpu.err
