#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_conv.py -x -q -k "edgenext" > gpurun_out/r2o_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2o_pytest.log
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/r2o_prof_update_b8.log 2>&1; grep -E "disp_head|TOTAL|back-to-back" gpurun_out/r2o_prof_update_b8.log
for v in 1 0 1 0; do
  FSB_FUSED_BLOCK=$v FSB_BENCH_WATCHDOG=60 timeout 90 python bench.py --steps 6 --warmup 3 --no-cpu > gpurun_out/r2o_bench_fb$v.json 2> gpurun_out/r2o_bench_fb$v.err
  echo "fused_block=$v rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/r2o_bench_fb$v.json | head -1) $(grep -o '"us_per_gru_iter": [0-9.]*' gpurun_out/r2o_bench_fb$v.json | head -1) $(grep -o '"sm_mhz": [0-9.]*' gpurun_out/r2o_bench_fb$v.json)"
done
