#!/bin/bash
# r3e: convd1 / convd2 on the side lane of the pipelined loop: parity + A/B
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "pipelined" > gpurun_out/r3e_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r3e_pytest.log
for v in 0 1 0 1; do
  FSB_ENCD_SIDE=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r3e_bench_encd$v.json 2> gpurun_out/r3e_bench_encd$v.err; echo "encd_side=$v rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r3e_bench_encd$v.json"))
print("encd_side=$v", round(d["value"],2), "pairs/s  e2e", round(d["e2e"]["value"],2), " us/iter", round(d["us_per_gru_iter"],1), d["clocks"]["sm_mhz"])
PY
done
