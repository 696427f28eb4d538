#!/bin/bash
# pass r2b: parity after the tail kernels (FFMA2 dwconv, one-MUFU GELU, dot head4, loop graph), bench, per-launch, ncu of the step
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2b_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"
FSB_LOOP_GRAPH=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2b_bench_itergraph.json 2> gpurun_out/r2b_bench_itergraph.err; echo "bench(per-iteration graph) rc=$?"
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/r2b_prof_update_b8.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv_umma|edgenext_mlp|dwconv|conv3x3_cout1' -s 23 -c 23 -o gpurun_out/prof_r2b_step -f python tools/prof_update.py 120 160 1 8 > gpurun_out/r2b_ncu_step.log 2>&1; tail -1 gpurun_out/r2b_ncu_step.log
ls -la gpurun_out | tail -12
