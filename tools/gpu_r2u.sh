#!/bin/bash
# r2u: throughput of the loop with the lookup fused into encoder.convc1 (A/B, twice each)
mkdir -p gpurun_out
for f in 0 1 0 1; do
  FSB_FUSED_LOOKUP_CONV=$f timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2u_bench_fused$f.json 2> gpurun_out/r2u_bench_fused$f.err; echo "fused=$f rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2u_bench_fused$f.json"))
print("fused=$f", round(d["value"],2), "pairs/s  e2e", round(d["e2e"]["value"],2), " us/iter", round(d["us_per_gru_iter"],1), "convc1", round(d["roofline"]["per_launch_ms"]["encoder.convc1"],4), d["clocks"]["sm_mhz"])
PY
done
