#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2n_pytest.log
for v in 0 1 0 1; do
  FSB_POOL_FP32=$v FSB_BENCH_WATCHDOG=60 timeout 90 python bench.py --steps 6 --warmup 3 --no-cpu > gpurun_out/r2n_bench_fp32pool$v.json 2> gpurun_out/r2n_bench_fp32pool$v.err
  echo "pool_fp32=$v rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/r2n_bench_fp32pool$v.json | head -1) $(grep -o '"us_per_gru_iter": [0-9.]*' gpurun_out/r2n_bench_fp32pool$v.json | head -1) $(grep -o '"sm_mhz": [0-9.]*' gpurun_out/r2n_bench_fp32pool$v.json)"
done
