#!/bin/bash
GNNX_GANG_LAUNCH=1 timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
for g in 0 1; do GNNX_GANG_LAUNCH=$g timeout 300 python bench.py --workload cfg2 --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('gang $g', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3))"; done
