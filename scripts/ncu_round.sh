#!/bin/bash
# usage: scripts/ncu_round.sh <tag>   (GPU box; writes gpurun_out/launches_<tag>_cfg2_tf32.csv and gpurun_out/ncu_<tag>_gemm_set_full.csv)
# 1) the un-profiled command must exit 0 first; 2) launch list: time + DRAM bytes per launch of 2 training steps;
# 3) one --set full capture of 12 consecutive tcgen05 GEMM launches of the second step.
TAG=${1:-r1g}
set -x
python scripts/profile_step.py --steps 2 > gpurun_out/profile_step_$TAG.log 2>&1 || { tail -5 gpurun_out/profile_step_$TAG.log; exit 1; }
ncu --profile-from-start off --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none \
    --csv --log-file gpurun_out/launches_${TAG}_cfg2_tf32.csv python scripts/profile_step.py --steps 2 > gpurun_out/ncu_list_$TAG.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:k_gemm_tf32 --launch-skip 60 --launch-count 12 \
    -o /tmp/ncu_${TAG}_gemm_set_full -f python scripts/profile_step.py --steps 1 > gpurun_out/ncu_full_$TAG.log 2>&1
ncu -i /tmp/ncu_${TAG}_gemm_set_full.ncu-rep --page raw --csv > gpurun_out/ncu_${TAG}_gemm_set_full_raw.csv 2>/dev/null
ls -la gpurun_out | tail -8
