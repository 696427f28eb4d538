#!/bin/bash
# helper kernels (7x7 one-channel conv, pool2x / interp, depthwise 7x7, init correlation): parity, then per-launch times
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for B in 4 8; do
  echo "== B=$B"
  timeout 300 python tools/prof_update.py 120 160 10 $B 2>&1 | grep -E "dwconv|convd1|pool|interp|mlp|TOTAL|back-to-back"
done
timeout 300 python tools/prof_kernels.py 8 2>&1 | tail -12
