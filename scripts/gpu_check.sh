#!/bin/bash
mkdir -p gpurun_out
python scripts/profile_step.py --steps 1 > /dev/null 2>&1 || exit 1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'k_bn_fwd_coop|k_bn_bwd_coop|k_gemm_tf32_chain|k_gemm_tf32_red' --launch-skip 8 --launch-count 14 \
    -o /tmp/ncu_r2k_new -f python scripts/profile_step.py --steps 1 > gpurun_out/ncu_r2k_new.log 2>&1
ncu -i /tmp/ncu_r2k_new.ncu-rep --page raw --csv > gpurun_out/ncu_r2k_new_raw.csv 2>/dev/null
python scripts/ncu_summarize.py gpurun_out/ncu_r2k_new_raw.csv "r2k new kernels" | head -90
