#!/bin/bash
# parity + bench + ncu of kernels 1-3 at 4 pairs
mkdir -p gpurun_out
bash tools/gpu_quick.sh
timeout 300 python tools/prof_kernels.py 3 4 > gpurun_out/prof_plain.log 2>&1; tail -1 gpurun_out/prof_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'volume_kernel|lookup_nhwc_kernel|lookup_nchw_kernel|corr_umma|geo_pyramid' -s 5 -c 5 -o gpurun_out/prof_r1g -f python tools/prof_kernels.py 3 4 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
