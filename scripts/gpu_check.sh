#!/bin/bash
# GPU box: the whole GPU suite, smoke(), then the round's bench lines (1 GPU) and in-situ traces.  Logs under gpurun_out/.
TAG=${1:-r2h}
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log; grep -E "^(FAILED|ERROR)" gpurun_out/pytest_gpu.log | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for wl in cfg2 cfg3 transformer cfg5; do
  python bench.py --workload $wl > gpurun_out/bench_${TAG}_${wl}_1gpu.json 2> gpurun_out/bench_${TAG}_${wl}.err; echo "$wl exit $?"; head -c 230 gpurun_out/bench_${TAG}_${wl}_1gpu.json; echo
done
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_reference_arm.json 2>/dev/null; head -c 300 gpurun_out/bench_${TAG}_reference_arm.json; echo
python scripts/kernel_trace.py --workload cfg2 --steps 2 --out gpurun_out/trace_${TAG}_cfg2_insitu.txt > /dev/null 2>&1
python scripts/kernel_trace.py --workload cfg3 --steps 2 --out gpurun_out/trace_${TAG}_cfg3_insitu.txt > /dev/null 2>&1
