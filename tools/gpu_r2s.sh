#!/bin/bash
# r2s: lookup fused into encoder.convc1 (csrc/lookconv.cu): parity, stand-alone timing, throughput in the loop
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "lookup_fused_into_convc1 or loop_with_lookup_fused" -s > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2s_pytest.log
timeout 300 python tools/lookconv_exp.py 8 2>&1 | tail -2
timeout 300 python tools/lookconv_exp.py 1 2>&1 | tail -1
if [ "$1" == "bench" ]; then
for f in 0 1 0 1; do
  FSB_FUSED_LOOKUP_CONV=$f timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2s_bench_fused$f.json 2> gpurun_out/r2s_bench_fused$f.err; echo "fused=$f rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2s_bench_fused$f.json"))
print("fused=$f", round(d["value"],2), "pairs/s  e2e", round(d["e2e"]["value"],2), " us/iter", round(d["us_per_gru_iter"],1), "convc1", d["roofline"]["per_launch_ms"]["encoder.convc1"], d["clocks"]["sm_mhz"])
PY
done
fi
