#!/bin/bash
# evidence pass r1m: parity, smoke, bench (both arms), launch list, ncu full of one update step and of kernels 1-3
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/prof_update_b8.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'conv_umma|conv_halo|edgenext_mlp|dwconv' -s 24 -c 24 -o gpurun_out/prof_r1m_step -f python tools/prof_update.py 120 160 1 8 > gpurun_out/ncu_step.log 2>&1; tail -1 gpurun_out/ncu_step.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'volume_kernel|lookup_nhwc_kernel|lookup_nchw_kernel|corr_tma|geo_pyramid|disp_regression|context_upsample' -s 7 -c 7 -o gpurun_out/prof_r1m_k123 -f python tools/prof_kernels.py 3 8 > gpurun_out/ncu_k123.log 2>&1; tail -1 gpurun_out/ncu_k123.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/launches_r1m.csv python bench.py --steps 1 --warmup 3 > gpurun_out/ncu_launches.log 2>&1; tail -1 gpurun_out/ncu_launches.log
ls -la gpurun_out | head -30
