#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/volstem_exp.py 8 > /dev/null 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:volume_stem -s 2 -c 1 -o gpurun_out/prof_r2x_volstem_f16 -f python tools/volstem_exp.py 8 > gpurun_out/r2x_ncu.log 2>&1; tail -1 gpurun_out/r2x_ncu.log
