#!/bin/bash
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log; grep -E "^(FAILED|ERROR)" gpurun_out/pytest_gpu.log | head
python scripts/kernel_trace.py --workload cfg2 --steps 2 --out gpurun_out/trace_r2e_cfg2_insitu.txt > /dev/null 2>&1; head -3 gpurun_out/trace_r2e_cfg2_insitu.txt; grep -E "embed_bwd|aggregate_bwd|reduce_partials" gpurun_out/trace_r2e_cfg2_insitu.txt
