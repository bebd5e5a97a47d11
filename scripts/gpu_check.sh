#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_collator.py tests/test_gpu_infer.py tests/test_dataset.py -m gpu -q -x 2>&1 | tail -5
timeout 600 python bench.py --workload cfg5 --full-pass 400000 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:d[k] for k in ('value','wall_s','store_upload_s','screen_s','value_without_upload','finite')})"
