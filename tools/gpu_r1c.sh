#!/bin/bash
# GPU pass r1c: parity tests, smoke, bench, per-launch update-block times, launch list, full ncu of conv kernels
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python tools/prof_update.py 120 160 20 > gpurun_out/prof_update.log 2>&1; echo "prof rc=$?"
cat gpurun_out/prof_update.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'conv_umma' -c 30 -o gpurun_out/prof_r1c_conv -f python tools/prof_update.py 120 160 1 > gpurun_out/ncu_conv.log 2>&1
tail -3 gpurun_out/ncu_conv.log
ls -la gpurun_out
