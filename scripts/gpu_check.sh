#!/bin/bash
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -2
for i in 1 2; do timeout 400 python bench.py --workload cfg2 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print(round(d['ms_per_step'],3), 'e2e', round(e['ms_per_step'],3), 'exact', round(e['exact_signature_ms_per_step'],3), d['gpu_launches'])"; done
