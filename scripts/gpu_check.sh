#!/bin/bash
# GPU box: round-2 bench lines at HEAD (1 GPU) + in-situ traces.  JSON under gpurun_out/.
for wl in cfg2 cfg3 transformer cfg5; do
  python bench.py --workload $wl > gpurun_out/bench_r2f_${wl}_1gpu.json 2> gpurun_out/bench_r2f_${wl}.err; echo "$wl exit $?"; head -c 260 gpurun_out/bench_r2f_${wl}_1gpu.json; echo
done
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_r2f_reference_arm.json 2>/dev/null; head -c 400 gpurun_out/bench_r2f_reference_arm.json; echo
python scripts/kernel_trace.py --workload cfg2 --steps 2 --out gpurun_out/trace_r2f_cfg2_insitu.txt > /dev/null 2>&1
python scripts/kernel_trace.py --workload cfg3 --steps 2 --out gpurun_out/trace_r2f_cfg3_insitu.txt > /dev/null 2>&1
python scripts/host_time.py cfg2 > gpurun_out/host_time_r2f.txt 2>&1; python scripts/host_time.py cfg3 >> gpurun_out/host_time_r2f.txt 2>&1; cat gpurun_out/host_time_r2f.txt
