#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q 2>&1 | tail -2
