#!/bin/bash
python -m pytest tests/test_gpu_gemm_tc.py tests/test_gpu_tf32_model.py tests/test_gpu_fullsize.py -m gpu -q 2>&1 | tail -4
for c in 2 4; do echo "GNNX_GEMM_CLUSTER=$c"; GNNX_GEMM_CLUSTER=$c python scripts/kernel_trace.py --workload cfg2 --steps 2 2>/dev/null | grep -E "^cfg2|k_gemm_tf32_rows2" ; done
