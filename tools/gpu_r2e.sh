#!/bin/bash
mkdir -p gpurun_out
for B in 12 16; do
  timeout 70 python tools/hang_probe.py $B 6 > gpurun_out/r2e_probe_sync_b$B.log 2>&1; echo "probe sync B=$B rc=$?"; tail -3 gpurun_out/r2e_probe_sync_b$B.log | cut -c1-300
  PROBE_MODE=lanes timeout 70 python tools/hang_probe.py $B 12 > gpurun_out/r2e_probe_lanes_b$B.log 2>&1; echo "probe lanes B=$B rc=$?"; tail -3 gpurun_out/r2e_probe_lanes_b$B.log | cut -c1-300
  FSB_TILE="gru04.zr=128" PROBE_MODE=lanes timeout 70 python tools/hang_probe.py $B 12 > gpurun_out/r2e_probe_lanes_zr128_b$B.log 2>&1; echo "probe lanes zr128 B=$B rc=$?"; tail -2 gpurun_out/r2e_probe_lanes_zr128_b$B.log | cut -c1-300
done
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "band" > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2e_pytest.log
