#!/bin/bash
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -2
for w in cfg2 cfg3; do
timeout 300 python scripts/kernel_trace.py --workload $w --steps 2 2>/dev/null | grep -E "^cfg|k_code_sum" | cut -c1-150
timeout 400 python bench.py --workload $w --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('$w', round(d['ms_per_step'],3), 'e2e', round(e['ms_per_step'],3), e['last_loss'])"
done
