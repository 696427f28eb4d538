#!/bin/bash
# evidence pass r3b at HEAD, part 1: smoke + bench, both arms as the driver runs them, and the other configurations
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r3b_bench_ref.json 2> gpurun_out/r3b_bench_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r3b_bench.json 2> gpurun_out/r3b_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --pairs 1 --no-cpu > gpurun_out/r3b_bench_pairs1.json 2> gpurun_out/r3b_bench_pairs1.err; echo "pairs1 rc=$?"
for cfgname in kitti64 vitl_hiera erp; do
  timeout 600 python bench.py --config $cfgname --steps 5 --warmup 3 --no-cpu > gpurun_out/r3b_bench_$cfgname.json 2> gpurun_out/r3b_bench_$cfgname.err; echo "$cfgname rc=$?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r3b_bench*.json")):
    try:
        d = json.load(open(f))
        print(f, round(d["value"], 3), d["unit"], "e2e", d.get("e2e", {}).get("value"), "frac", d.get("roofline", {}).get("frac"), d.get("clocks"))
    except Exception as e:
        print(f, "unreadable", e)
PY
