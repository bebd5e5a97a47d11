#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_gemm_pair.py -m gpu -q -x 2>&1 | tail -3
SWEEP=0,1 timeout 600 python scripts/gemm_shapes.py 2>&1 | tail -12
