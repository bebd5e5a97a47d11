#!/bin/bash
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout 300 python scripts/kernel_trace.py --workload cfg2 --steps 2 2>/dev/null | grep -E "^cfg2|k_gemm_tf32" | cut -c1-150
timeout 400 python bench.py --workload cfg2 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['roofline']['frac'])"
