#!/bin/bash
# r2r: pipelined loop (coarse GRU levels of iteration i+1 under the disparity head of iteration i): parity + throughput
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipelined or fused_loop" > gpurun_out/r2r_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2r_pytest.log
for p in 0 1 0 1; do
  FSB_PIPELINE=$p timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2r_bench_pipe$p.json 2> gpurun_out/r2r_bench_pipe$p.err; echo "pipe=$p rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2r_bench_pipe$p.json"))
print("pipe=$p", round(d["value"],2), "pairs/s  e2e", round(d["e2e"]["value"],2), " us/iter", round(d["us_per_gru_iter"],1), d["clocks"])
PY
done
