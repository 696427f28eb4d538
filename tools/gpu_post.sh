#!/bin/bash
# post kernels (disparity regression, convex upsampling): parity, per-kernel times in the bench line, ncu full
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'disp_regression|context_upsample|corr_umma' -s 6 -c 6 -o gpurun_out/prof_post -f python tools/prof_kernels.py 4 8 > gpurun_out/ncu_post.log 2>&1; tail -1 gpurun_out/ncu_post.log
