#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_ops.py tests/test_zz_no_alloc.py tests/test_gpu_model.py -m gpu -q -x 2>&1 | tail -3
timeout 300 python scripts/kernel_trace.py --workload cfg3 --steps 2 2>/dev/null | grep -E "^cfg|k_colsum" | cut -c1-150
timeout 400 python bench.py --workload cfg3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('cfg3', round(d['ms_per_step'],3), 'e2e', round(e['ms_per_step'],3))"
