#!/bin/bash
mkdir -p gpurun_out
for w in cfg2 cfg3 transformer cfg5; do
  timeout 600 python bench.py --workload $w > gpurun_out/bench_r2k_${w}_1gpu.json 2> gpurun_out/bench_r2k_$w.err; tail -c 300 gpurun_out/bench_r2k_${w}_1gpu.json; echo
done
timeout 600 python bench.py --impl reference > gpurun_out/bench_r2k_reference_arm.json 2> gpurun_out/bench_r2k_ref.err; tail -c 400 gpurun_out/bench_r2k_reference_arm.json
for w in cfg2 cfg3; do timeout 300 python scripts/kernel_trace.py --workload $w --steps 2 > gpurun_out/trace_r2k_${w}_insitu.txt 2>/dev/null; head -3 gpurun_out/trace_r2k_${w}_insitu.txt | cut -c1-200; done
