#!/bin/bash
# evidence pass r2y at HEAD: full parity, smoke, bench both arms as the driver runs them, other configs, per-launch, launch list, ncu full
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2y_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r2y_bench_ref.json 2> gpurun_out/r2y_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --pairs 1 --no-cpu > gpurun_out/r2y_bench_pairs1.json 2> gpurun_out/r2y_bench_pairs1.err; echo "pairs1 rc=$?"
for cfgname in kitti64 vitl_hiera erp; do
  timeout 600 python bench.py --config $cfgname --steps 5 --warmup 3 --no-cpu > gpurun_out/r2y_bench_$cfgname.json 2> gpurun_out/r2y_bench_$cfgname.err; echo "$cfgname rc=$?"
done
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/r2y_prof_update_b8.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv_umma|edgenext_mlp|dwconv|conv3x3_cout1' -s 23 -c 23 -o gpurun_out/prof_r2y_step -f python tools/prof_update.py 120 160 1 8 > gpurun_out/r2y_ncu_step.log 2>&1; tail -1 gpurun_out/r2y_ncu_step.log
timeout 300 python tools/prof_kernels.py 3 8 > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'volume_kernel|lookup_nhwc_kernel|lookup_nchw_kernel|corr_tma|geo_pyramid|disp_regression|context_upsample' -s 9 -c 9 -o gpurun_out/prof_r2y_k123 -f python tools/prof_kernels.py 3 8 > gpurun_out/r2y_ncu_k123.log 2>&1; tail -1 gpurun_out/r2y_ncu_k123.log
timeout 300 python tools/lookconv_exp.py 8 > gpurun_out/r2y_lookconv.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lookup_convc1 -s 3 -c 1 -o gpurun_out/prof_r2y_lookconv -f python tools/lookconv_exp.py 8 > gpurun_out/r2y_ncu_lookconv.log 2>&1; tail -1 gpurun_out/r2y_ncu_lookconv.log
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/r2y_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/r2y_ncu_launches.log 2>&1; tail -1 gpurun_out/r2y_ncu_launches.log
