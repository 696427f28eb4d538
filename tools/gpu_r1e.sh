#!/bin/bash
# GPU pass r1e: parity tests, smoke, bench, per-launch update-block times, launch list, full ncu of the conv kernel
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 300 python tools/prof_update.py 120 160 20 4 > gpurun_out/prof_update_b4.log 2>&1; echo "prof rc=$?"
cat gpurun_out/prof_update_b4.log
timeout 300 python tools/prof_update.py 120 160 20 1 > gpurun_out/prof_update_b1.log 2>&1; echo "prof rc=$?"
tail -3 gpurun_out/prof_update_b1.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 48 -c 24 -o gpurun_out/prof_r1e_conv -f python tools/prof_update.py 120 160 1 4 > gpurun_out/ncu_conv.log 2>&1
tail -3 gpurun_out/ncu_conv.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 400 --csv --log-file gpurun_out/launches_r1e.csv python bench.py --steps 1 --warmup 3 --pairs 4 > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/ncu_launches.log
ls -la gpurun_out
