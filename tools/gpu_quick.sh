#!/bin/bash
# quick GPU pass: parity tests, smoke, bench at 4 pairs
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --warmup 3 --pairs ${PAIRS:-4} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench.json'))
print('value',round(d['value'],2),'ms/step',round(d['ms_per_step'],2),'us/iter',round(d['us_per_gru_iter'],1),'e2e',round(d['e2e']['value'],2),'conv TF',round(d['roofline']['achieved'],1),'launches',d['gpu_launches'])
for k,v in d['roofline']['hbm_kernels'].items(): print(f"  {k:32s} {v['ms']*1e3:8.1f} us  {v['achieved_gbs']:7.0f} GB/s  {v['frac_of_measured_hbm']:.3f}")
PY
