#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_gatv2_h2.py -m gpu -q 2>&1 | grep -E "^E  |FAILED|passed|failed" | head -20
