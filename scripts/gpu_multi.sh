#!/bin/bash
# usage: scripts/gpu_multi.sh <N> <tag>  (GPU box with N GPUs): NCCL parity test (N >= 2), weak and strong scaling lines, and at
# N == 8 the full-size cfg5 screening pass.  JSON lines under gpurun_out/.
N=$1; TAG=${2:-r2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
if [ "$N" -ge 2 ]; then python -m pytest tests/test_gpu_dp_nccl.py -m gpu -q 2>&1 | tail -3; fi
$TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_${TAG}_cfg4_weak_${N}gpu.json 2> gpurun_out/bench_${TAG}_weak_${N}.err; echo "weak exit $?"; head -c 330 gpurun_out/bench_${TAG}_cfg4_weak_${N}gpu.json; echo
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --scaling strong > gpurun_out/bench_${TAG}_cfg4_strong_${N}gpu.json 2> gpurun_out/bench_${TAG}_strong_${N}.err; echo "strong exit $?"; head -c 330 gpurun_out/bench_${TAG}_cfg4_strong_${N}gpu.json; echo
if [ "$N" -eq 8 ]; then
  $TR bench.py --gpus $N --workload cfg5 --full-pass 1250000 > gpurun_out/bench_${TAG}_cfg5_10M_8gpu.json 2> gpurun_out/bench_${TAG}_cfg5_full.err; echo "cfg5 full exit $?"; head -c 700 gpurun_out/bench_${TAG}_cfg5_10M_8gpu.json; echo; tail -3 gpurun_out/bench_${TAG}_cfg5_full.err
  $TR bench.py --gpus $N --workload cfg5 --steps 20 --no-cpu-baseline > gpurun_out/bench_${TAG}_cfg5_infer_8gpu.json 2>/dev/null; head -c 300 gpurun_out/bench_${TAG}_cfg5_infer_8gpu.json; echo
fi
tail -n 3 gpurun_out/bench_${TAG}_weak_${N}.err; tail -n 3 gpurun_out/bench_${TAG}_strong_${N}.err
