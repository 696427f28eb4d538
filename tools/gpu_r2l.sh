#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_configs.py -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2l_pytest.log
for v in 1 0 1 0; do
  FSB_TAIL_POOLS=$v FSB_BENCH_WATCHDOG=60 timeout 90 python bench.py --steps 6 --warmup 3 --no-cpu > gpurun_out/r2l_bench_tp$v.json 2> gpurun_out/r2l_bench_tp$v.err
  echo "tail_pools=$v rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/r2l_bench_tp$v.json | head -1) $(grep -o '"us_per_gru_iter": [0-9.]*' gpurun_out/r2l_bench_tp$v.json | head -1) $(grep -o '"sm_mhz": [0-9.]*' gpurun_out/r2l_bench_tp$v.json)"
done
