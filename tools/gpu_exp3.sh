#!/bin/bash
mkdir -p gpurun_out
SWEEP_ONLY="1x1" timeout 300 python tools/conv_sweep.py
SWEEP_ONLY="1x1" SWEEP_REPS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_umma -o gpurun_out/prof_1x1 -f python tools/conv_sweep.py > gpurun_out/ncu_1x1.log 2>&1
tail -2 gpurun_out/ncu_1x1.log
