#!/bin/bash
# r3f: final state check at HEAD: GPU suite twice, smoke, bench with default flags
mkdir -p gpurun_out
for i in 1 2; do
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r3f_pytest_$i.log 2>&1; echo "pytest $i rc=$?"; tail -1 gpurun_out/r3f_pytest_$i.log
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r3f_bench_default.json 2> gpurun_out/r3f_bench_default.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/r3f_bench_default.json"))
print(round(d["value"], 2), d["unit"], "steps", d["steps"], "warmup", d["warmup"], "e2e", round(d["e2e"]["value"], 2), "frac", round(d["roofline"]["frac"], 3),
      "cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["kind"], d["clocks"], "launches", d["gpu_launches"])
PY
