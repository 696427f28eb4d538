#!/bin/bash
# ncu full capture of the volume and lookup kernels at 4 pairs
mkdir -p gpurun_out
timeout 300 python tools/prof_kernels.py 3 4 > gpurun_out/prof_plain.log 2>&1; tail -2 gpurun_out/prof_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'volume_kernel|lookup_nhwc_kernel|lookup_kernel|corr_umma' -s 4 -c 4 -o gpurun_out/prof_r1f -f python tools/prof_kernels.py 3 4 > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
