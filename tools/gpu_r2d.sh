#!/bin/bash
# pass r2d: band-form tests, where does --pairs 16 hang (watchdog), memcheck over the small case
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_conv.py tests/test_abi.py -x -q -k "band or cout1 or head4 or abi or lookup or geo" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2d_pytest.log
FSB_BENCH_WATCHDOG=50 timeout 90 python bench.py --steps 2 --warmup 3 --pairs 16 --no-cpu > gpurun_out/r2d_bench_p16.json 2> gpurun_out/r2d_bench_p16.err; echo "bench pairs=16 rc=$?"; tail -40 gpurun_out/r2d_bench_p16.err
FSB_BENCH_WATCHDOG=50 timeout 90 python bench.py --steps 2 --warmup 3 --pairs 12 --no-cpu > gpurun_out/r2d_bench_p12.json 2> gpurun_out/r2d_bench_p12.err; echo "bench pairs=12 rc=$?"; tail -5 gpurun_out/r2d_bench_p12.err
bash tools/sanitize.sh memcheck
