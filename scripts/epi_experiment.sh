for m in 0 1 2 3; do echo "EPI $m"; GNNX_EPI_DEBUG=$m timeout 200 python scripts/bench_gemm.py 2>&1 | grep -E "^NT  " | head -3; done
