#!/bin/bash
# quick GPU pass: parity tests, per-launch update-block profile at batch 1 and 4, bench at 1/2/4 pairs
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/prof_update.py 120 160 10 4 2>&1 | tail -36
for p in 1 2 4; do
timeout 600 python bench.py --steps 3 --warmup 3 --pairs $p > gpurun_out/bench_p$p.json 2> gpurun_out/bench_p$p.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/bench_p$p.json'))
print('pairs',$p,'value',round(d['value'],2),'ms/step',round(d['ms_per_step'],2),'us/iter',round(d['us_per_gru_iter'],1),'e2e',round(d['e2e']['value'],2),'conv TF',round(d['roofline']['achieved'],1))
PY
done
