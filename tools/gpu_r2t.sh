#!/bin/bash
# r2t: fused lookup -> convc1 kernel: parity, timing, ncu --set full (one launch, 8 pairs)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lookup_fused_into_convc1 or loop_with_lookup_fused" -s > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2s_pytest.log
timeout 300 python tools/lookconv_exp.py 8 2>&1 | tail -1
timeout 300 python tools/lookconv_exp.py 1 2>&1 | tail -1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lookup_convc1 -s 3 -c 1 -o gpurun_out/prof_r2t_lookconv -f python tools/lookconv_exp.py 8 > gpurun_out/r2t_ncu.log 2>&1; tail -1 gpurun_out/r2t_ncu.log
