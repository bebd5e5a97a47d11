#!/bin/bash
mkdir -p gpurun_out
for w in cfg2 cfg3 transformer cfg5; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_r2l_${w}_1gpu.json 2> gpurun_out/bench_r2l_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r2l_${w}_1gpu.json').read().strip().splitlines()[-1])
print('$w', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), d['gpu_launches'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])
PY
done
for w in cfg2 cfg3; do timeout 300 python scripts/kernel_trace.py --workload $w --steps 2 > gpurun_out/trace_r2l_${w}_insitu.txt 2>/dev/null; head -1 gpurun_out/trace_r2l_${w}_insitu.txt | cut -c1-200; done
