#!/bin/bash
# pass r2c: new conv3x3_cout1 tiling + MLP epilogue; batch-size sweep (wave quantisation); other configs
mkdir -p gpurun_out
free -g | head -2; nproc
timeout 900 python -m pytest tests/test_gpu_conv.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c_pytest.log
for P in 8 16 24; do
  timeout 600 python bench.py --steps 6 --warmup 3 --pairs $P --no-cpu > gpurun_out/r2c_bench_p$P.json 2> gpurun_out/r2c_bench_p$P.err; echo "bench pairs=$P rc=$?"
done
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/r2c_prof_update_b8.log 2>&1; tail -8 gpurun_out/r2c_prof_update_b8.log
for C in vitl_hiera kitti64 erp; do
  timeout 900 python bench.py --steps 3 --warmup 3 --config $C --no-cpu > gpurun_out/r2c_bench_$C.json 2> gpurun_out/r2c_bench_$C.err; echo "bench $C rc=$?"; tail -c 300 gpurun_out/r2c_bench_$C.err
done
