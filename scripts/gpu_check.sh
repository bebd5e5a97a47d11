#!/bin/bash
mkdir -p gpurun_out
python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -2
for w in cfg2 cfg3 transformer; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_r2m_${w}_1gpu.json 2> gpurun_out/bench_r2m_$w.err; python - <<PY
import json
d=json.loads(open('gpurun_out/bench_r2m_${w}_1gpu.json').read().strip().splitlines()[-1])
e=d['e2e']
print('$w', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(e['value']), round(e['ms_per_step'],3), e.get('sync_every_step_ms_per_step'), e.get('exact_signature_ms_per_step'), d['gpu_launches'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])
PY
done
for w in cfg2 cfg3; do timeout 300 python scripts/kernel_trace.py --workload $w --steps 2 > gpurun_out/trace_r2m_${w}_insitu.txt 2>/dev/null; head -1 gpurun_out/trace_r2m_${w}_insitu.txt | cut -c1-160; done
