#!/bin/bash
mkdir -p gpurun_out
bash tools/gpu_quick.sh
echo "=== baseline tiles (auto)"; timeout 300 python tools/prof_update.py 120 160 10 4 2>&1 | grep -E "gru04|gru08\.(zr|q)|TOTAL|back-to-back"
echo "=== halo q,zr BN128"; FSB_HALO=gru04.conv0,gru04.conv1,gru04.q,gru04.zr FSB_BN=gru04.conv0=192,gru04.conv1=256,gru04.q=128,gru04.zr=128,encoder.convc2=256 timeout 300 python tools/prof_update.py 120 160 10 4 2>&1 | grep -E "gru04|gru08\.(zr|q)|convc2|TOTAL|back-to-back"
echo "=== tap q BN128"; FSB_HALO=gru04.conv0,gru04.conv1 FSB_BN=gru04.conv0=192,gru04.conv1=256,gru04.q=128,gru08.q=64,encoder.convc2=256 timeout 300 python tools/prof_update.py 120 160 10 4 2>&1 | grep -E "gru04|gru08\.(zr|q)|TOTAL|back-to-back"
echo "=== halo q,zr + gru08 halo"; FSB_HALO=gru04.conv0,gru04.conv1,gru04.q,gru04.zr,gru08.conv0,gru08.conv1,gru08.q,gru08.zr FSB_BN=gru04.conv0=192,gru04.conv1=256,gru04.q=128,gru04.zr=128,encoder.convc2=256,gru08.q=64,gru08.zr=64,gru08.conv0=128,gru08.conv1=128 timeout 300 python tools/prof_update.py 120 160 10 4 2>&1 | grep -E "gru04|gru08|TOTAL|back-to-back"
