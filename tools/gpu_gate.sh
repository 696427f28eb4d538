#!/bin/bash
# gate-epilogue TMA stores: parity, then A/B of the update step with FSB_GATE_STORE=0/1
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for gs in 0 1; do
  for B in 4 8; do
    echo "== FSB_GATE_STORE=$gs B=$B"
    FSB_GATE_STORE=$gs timeout 300 python tools/prof_update.py 120 160 10 $B 2>&1 | grep -E "zr|\.q |TOTAL|back-to-back"
  done
done
