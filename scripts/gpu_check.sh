#!/bin/bash
mkdir -p gpurun_out
python scripts/profile_step.py --workload cfg3 --steps 1 > /dev/null 2>&1 || exit 1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'k_gatv2_fwd_h2|k_gatv2_bwd_dst_h2|k_code_sum_sorted|k_code_sum_fold|k_colsum_coop4' --launch-skip 4 --launch-count 12 \
    -o /tmp/ncu_r2o_cfg3 -f python scripts/profile_step.py --workload cfg3 --steps 1 > gpurun_out/ncu_r2o_cfg3.log 2>&1
ncu -i /tmp/ncu_r2o_cfg3.ncu-rep --page raw --csv > gpurun_out/ncu_r2o_cfg3_raw.csv 2>/dev/null
python scripts/ncu_summarize.py gpurun_out/ncu_r2o_cfg3_raw.csv "r2o cfg3 new kernels" | grep -E "^----|duration|dram throughput|issue active|dram GB" | head -40
