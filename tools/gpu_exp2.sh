#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python tools/conv_sweep.py 2>&1 | grep "lay=0"
timeout 300 python tools/prof_update.py 120 160 10 4 2>&1 | tee gpurun_out/prof_update_b4.log
timeout 300 python tools/prof_update.py 120 160 10 1 2>&1 | tail -3
