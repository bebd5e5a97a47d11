#!/bin/bash
mkdir -p gpurun_out
python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -2
for w in cfg3 transformer; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_r2o_${w}_1gpu.json 2> gpurun_out/bench_r2o_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r2o_${w}_1gpu.json').read().strip().splitlines()[-1])
e=d['e2e']
print('$w', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(e['value']), round(e['ms_per_step'],3), d['gpu_launches'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])
PY
done
