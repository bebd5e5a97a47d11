#!/bin/bash
python -m pytest tests/test_gpu_ops.py tests/test_gpu_model.py -m gpu -q 2>&1 | tail -2
for b in 0 1 0 1; do echo "GNNX_GATHER8=$b"; GNNX_GATHER8=$b python scripts/kernel_trace.py --workload cfg2 --steps 2 2>/dev/null | grep -E "k_pna_gather_fwd" | cut -c1-150; done
