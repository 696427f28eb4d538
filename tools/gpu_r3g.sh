#!/bin/bash
# r3g: the GPU suite repeatedly after restricting the fused lookup -> convc1 to the pipelined order (bounded: a stall costs 200 s)
mkdir -p gpurun_out
for i in 1 2; do
  timeout 200 python -m pytest tests -m gpu -x -q > gpurun_out/r3g_pytest_$i.log 2>&1; echo "pytest $i rc=$?"; tail -1 gpurun_out/r3g_pytest_$i.log | cut -c1-120
done
timeout 200 python bench.py --steps 6 --warmup 3 --no-cpu > gpurun_out/r3g_bench.json 2> gpurun_out/r3g_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r3g_bench.json')); print(round(d['value'],2), round(d['e2e']['value'],2), round(d['roofline']['frac'],3), d['roofline']['hbm_kernels'].get('lookup_convc1',{}).get('ms'), d['clocks']['sm_mhz'])"
