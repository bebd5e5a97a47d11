#!/bin/bash
python -m pytest tests/test_gpu_model.py tests/test_gpu_ops.py tests/test_gpu_tf32_model.py tests/test_gpu_fullsize.py -m gpu -q 2>&1 | tail -3
python scripts/kernel_trace.py --workload cfg3 --steps 2 2>/dev/null | grep -E "^cfg3|gatv2|colsum" | cut -c1-150
