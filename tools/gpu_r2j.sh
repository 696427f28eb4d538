#!/bin/bash
# evidence pass r2j: full parity, smoke, memcheck, bench (both arms), per-launch, launch list, ncu full of the update step
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2j_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
bash tools/sanitize.sh memcheck
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2j_bench_ref.json 2> gpurun_out/r2j_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --steps 3 --warmup 3 --config erp --no-cpu > gpurun_out/r2j_bench_erp.json 2> gpurun_out/r2j_bench_erp.err; echo "erp rc=$?"
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/r2j_prof_update_b8.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv_umma|edgenext_mlp|dwconv|conv3x3_cout1' -s 23 -c 23 -o gpurun_out/prof_r2j_step -f python tools/prof_update.py 120 160 1 8 > gpurun_out/r2j_ncu_step.log 2>&1; tail -1 gpurun_out/r2j_ncu_step.log
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu > /dev/null 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 300 -c 400 --csv --log-file gpurun_out/r2j_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu > gpurun_out/r2j_ncu_launches.log 2>&1; tail -1 gpurun_out/r2j_ncu_launches.log
