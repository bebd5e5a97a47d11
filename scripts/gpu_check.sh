#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_gatv2_h2.py -m gpu -q -x 2>&1 | tail -6
timeout 600 python -m pytest tests/test_gpu_model.py -m gpu -q -x -k "gat or GAT" 2>&1 | tail -2
for p in 0 1; do echo "GNNX_GATV2_H2=$p"; GNNX_GATV2_H2=$p timeout 300 python scripts/kernel_trace.py --workload cfg3 --steps 2 2>/dev/null | grep -E "^cfg3|k_gatv2_fwd" | cut -c1-150; done
