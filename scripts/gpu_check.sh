#!/bin/bash
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log; grep -E "^(FAILED|ERROR)|Error" gpurun_out/pytest_gpu.log | head
python bench.py --steps 30 --no-cpu-baseline > gpurun_out/bench_r2g_cfg2.json 2> gpurun_out/bench_r2g.err; echo "exit $?"; tail -3 gpurun_out/bench_r2g.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/bench_r2g_cfg2.json") if l.startswith("{")][0])
print("value ms", d["ms_per_step"], "e2e", {k: v for k, v in d["e2e"].items() if k != "api"})
PY
