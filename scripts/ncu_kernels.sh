#!/bin/bash
# usage: scripts/ncu_kernels.sh <tag> [workload]   (GPU box) -- ncu --set full of every NON-GEMM gnnx kernel of one training
# step (gather / aggregate / BatchNorm / embedding / pooling / attention kernels): writes the report to /tmp (reports are tens of MB; gpurun_out/ is capped at 64 MiB)
# and its raw CSV to gpurun_out/; summarise here with scripts/ncu_summarize.py.  The un-profiled command must have exited 0 before.
TAG=${1:-r2a}; WL=${2:-cfg2}
set -x
python scripts/profile_step.py --workload $WL --steps 1 > gpurun_out/profile_step_${TAG}_${WL}.log 2>&1 || { tail -5 gpurun_out/profile_step_${TAG}_${WL}.log; exit 1; }
ncu --profile-from-start off --set full --clock-control none --import-source on \
    -k regex:'k_pna|k_bn_|k_embed|k_code|k_pool|k_gather_rows|k_gatv2|k_tconv|k_sum_partials|k_relhuber|k_adamw|k_permute' \
    --launch-count 160 -o /tmp/ncu_${TAG}_${WL}_kernels -f python scripts/profile_step.py --workload $WL --steps 1 > gpurun_out/ncu_kernels_${TAG}_${WL}.log 2>&1
ncu -i /tmp/ncu_${TAG}_${WL}_kernels.ncu-rep --page raw --csv > gpurun_out/ncu_${TAG}_${WL}_kernels_raw.csv 2>/dev/null
ls -la gpurun_out | tail -5
