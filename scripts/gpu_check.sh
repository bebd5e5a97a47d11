#!/bin/bash
GNNX_GEMM_EPI8=1 python -m pytest tests/test_gpu_gemm_tc.py tests/test_gpu_degree_groups.py tests/test_gpu_tf32_model.py -m gpu -q 2>&1 | tail -2
for b in 0 1 0 1; do echo "GNNX_GEMM_EPI8=$b"; GNNX_GEMM_EPI8=$b python scripts/kernel_trace.py --workload cfg2 --steps 2 2>/dev/null | grep -E "^cfg2|k_gemm_tf32_rows2" | cut -c1-150; done
