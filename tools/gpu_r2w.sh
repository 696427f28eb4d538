#!/bin/bash
# r2w: cost volume fused with corr_stem[0] (csrc/volstem.cu): parity, install(fused_stem=True) through the unchanged caller, timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "volume_fused_with_corr_stem or volume_stem_rejects" > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2w_pytest.log
timeout 900 python -m pytest tests/test_gpu_e2e_model.py -m gpu -x -q -s -k "install_matches_stock_forward or uninstall" > gpurun_out/r2w_e2e.log 2>&1; echo "e2e rc=$?"; grep "\[e2e\]" gpurun_out/r2w_e2e.log; tail -2 gpurun_out/r2w_e2e.log
timeout 300 python tools/volstem_exp.py 8 2>&1 | tail -2
