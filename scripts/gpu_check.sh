#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -4
for w in cfg2 cfg3 transformer; do
  timeout 400 python bench.py --workload $w > gpurun_out/bench_r2i_$w.json 2> gpurun_out/bench_r2i_$w.err; tail -c 1500 gpurun_out/bench_r2i_$w.json | cut -c1-700
done
