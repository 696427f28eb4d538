#!/bin/bash
# first GPU pass: parity tests, smoke, bench, launch list, full ncu capture of kernels 1-3
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -5 gpurun_out/smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python tools/prof_kernels.py 3 > gpurun_out/prof_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'volume_kernel|lookup_kernel|geo_pyramid_kernel|corr_fp32_kernel' -s 4 -c 8 -o gpurun_out/prof_r1a -f python tools/prof_kernels.py 3 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 400 --csv --log-file gpurun_out/launches_r1a.csv python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/ncu_launches.log
ls -la gpurun_out
