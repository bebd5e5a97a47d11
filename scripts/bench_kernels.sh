#!/bin/bash
# usage: scripts/bench_kernels.sh <workload> [steps]  -- prints step time and the per-entry-point table of a bench run
W=${1:-cfg2}; S=${2:-10}
timeout 400 python bench.py --workload $W --steps $S --warmup 3 --no-cpu-baseline > gpurun_out/bench_$W.json 2> gpurun_out/bench_$W.err || tail -n 5 gpurun_out/bench_$W.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$W.json").read().strip().splitlines()[-1])
print("$W", round(d["value"]), "graphs/s", round(d["ms_per_step"],3), "ms/step  e2e", round(d["e2e"]["value"]))
for k,v in d["kernels"].items(): print(f"  {k:28s} x{v['calls']:4d} {v['ms_per_call']:.4f} ms  {v['frac_of_hbm_peak']:.3f} of HBM peak  share {v['share_of_gnnx_time']:.3f}")
PY
