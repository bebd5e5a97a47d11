#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "batchnorm" 2>&1 | grep "AssertionError\|passed\|failed" | head -5
GNNX_BN_COOP=0 timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "batchnorm" 2>&1 | tail -1
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -2
