#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_conv.py tests/test_gpu_parity.py -x -q -s -k "cout1 or head4 or midsize or loop" > gpurun_out/r2k_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|update block errors" gpurun_out/r2k_pytest.log | cut -c1-900
timeout 300 python tools/prof_update.py 120 160 10 8 > gpurun_out/r2k_prof_update_b8.log 2>&1; grep -E "conv.4|mlp|dwconv|TOTAL|back-to-back" gpurun_out/r2k_prof_update_b8.log
bash tools/sanitize.sh memcheck
