#!/bin/bash
# GPU box: the whole GPU suite, smoke(), a short bench.  Logs (small) under gpurun_out/.
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log; grep -E "^(FAILED|ERROR)" gpurun_out/pytest_gpu.log | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --steps 20 > gpurun_out/bench_r2d_cfg2.json 2> gpurun_out/bench_r2d_cfg2.err; echo "bench exit $?"; head -c 300 gpurun_out/bench_r2d_cfg2.json; echo; tail -2 gpurun_out/bench_r2d_cfg2.err
