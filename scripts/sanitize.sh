#!/bin/bash
# usage: scripts/sanitize.sh <tag>   (GPU box) -- compute-sanitizer over the small-shape GPU tests (SURVEY section 5 "race detection"):
# memcheck over the per-kernel suites, racecheck (shared-memory hazards) and synccheck over the tcgen05 / mbarrier kernels'
# tests.  Summaries go to gpurun_out/sanitize_<tag>.log; copy it to profiles/.
TAG=${1:-r2}
LOG=gpurun_out/sanitize_$TAG.log
: > $LOG
run() {   # tool, pytest selection...
    local tool=$1; shift
    echo "==== compute-sanitizer --tool $tool : pytest $*" >> $LOG
    timeout 1500 compute-sanitizer --tool $tool --print-limit 5 --error-exitcode 0 python -m pytest "$@" -m gpu -q -x -p no:cacheprovider > /tmp/san.out 2>&1
    grep -E "^=========|passed|failed|error" /tmp/san.out | grep -vE "^=========\s*$" | tail -25 >> $LOG
}
run memcheck tests/test_gpu_ops.py tests/test_gpu_collator.py tests/test_gpu_degree_groups.py tests/test_zz_no_alloc.py
run memcheck tests/test_gpu_gemm_tc.py -k "not 200000 and not 100000"
run memcheck tests/test_gpu_pna_fused.py -k "700-4-64 or 1200-9-64 or 900-0-64"
run racecheck tests/test_gpu_gemm_tc.py -k "not 200000 and not 100000"
run racecheck tests/test_gpu_pna_fused.py -k "700-4-64"
run synccheck tests/test_gpu_gemm_tc.py -k "not 200000 and not 100000"
run memcheck tests/test_gpu_model.py -k "PNA or GATv2 or Transformer"
tail -60 $LOG
