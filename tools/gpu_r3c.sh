#!/bin/bash
# evidence pass r3c at HEAD, part 2: per-launch times, ncu --set full tables (converted on the box: the captures themselves
# exceed what gpurun copies back), launch list of the bench
mkdir -p gpurun_out
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/r3c_update_block_per_launch_b8.txt 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv_umma|edgenext_mlp|dwconv|conv3x3_cout1' -s 23 -c 23 -o /tmp/prof_step -f python tools/prof_update.py 120 160 1 8 > gpurun_out/r3c_ncu_step.log 2>&1; tail -1 gpurun_out/r3c_ncu_step.log
python tools/ncu_table.py /tmp/prof_step.ncu-rep > gpurun_out/r3c_ncu_full_update_step.txt 2>&1
timeout 300 python tools/prof_kernels.py 3 8 > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'volume_kernel|lookup_nhwc_kernel|lookup_nchw_kernel|corr_tma|geo_pyramid|disp_regression|context_upsample' -s 9 -c 9 -o /tmp/prof_k123 -f python tools/prof_kernels.py 3 8 > gpurun_out/r3c_ncu_k123.log 2>&1; tail -1 gpurun_out/r3c_ncu_k123.log
python tools/ncu_table.py /tmp/prof_k123.ncu-rep > gpurun_out/r3c_ncu_full_kernels_1_2_2b_3_post.txt 2>&1
timeout 300 python tools/lookconv_exp.py 8 > gpurun_out/r3c_lookconv.txt 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lookup_convc1 -s 3 -c 1 -o /tmp/prof_lookconv -f python tools/lookconv_exp.py 8 > gpurun_out/r3c_ncu_lookconv.log 2>&1; tail -1 gpurun_out/r3c_ncu_lookconv.log
python tools/ncu_table.py /tmp/prof_lookconv.ncu-rep > gpurun_out/r3c_ncu_full_lookup_convc1.txt 2>&1
python tools/ncu_lines.py /tmp/prof_lookconv.ncu-rep lookup_convc1 16 >> gpurun_out/r3c_ncu_full_lookup_convc1.txt 2>&1
python tools/make_traffic.py /tmp/prof_step.ncu-rep /tmp/prof_k123.ncu-rep /tmp/prof_lookconv.ncu-rep > gpurun_out/r3c_traffic.json 2> gpurun_out/r3c_traffic.err
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/r3c_launches_bench.csv python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/r3c_ncu_launches.log 2>&1; tail -c 300 gpurun_out/r3c_ncu_launches.log
ls -la gpurun_out | tail -15
