#!/bin/bash
# row-reuse mode of the pair-mode conv kernel: parity, then A/B of the update step with FSB_CONV_VR=0/1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
for vr in 0 1; do
  for B in 4 8; do
    echo "== FSB_CONV_VR=$vr B=$B"
    FSB_CONV_VR=$vr timeout 300 python tools/prof_update.py 120 160 10 $B 2>&1 | grep -E "TFLOP|back-to-back" | grep -v -E "mlp"
  done
done
