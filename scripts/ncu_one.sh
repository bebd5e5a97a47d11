#!/bin/bash
# usage: scripts/ncu_one.sh <kernel regex> <tag> [skip] [count]  (GPU box): ncu --set full of a few launches of one kernel
# inside a cfg2 training step; prints the summary, the stall-reason columns and the hottest SASS lines.
K=$1; TAG=$2; SKIP=${3:-2}; CNT=${4:-1}
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:$K --launch-skip $SKIP --launch-count $CNT \
    -o /tmp/ncu_$TAG -f python scripts/profile_step.py --steps 1 > gpurun_out/ncu_one_$TAG.log 2>&1
ncu -i /tmp/ncu_$TAG.ncu-rep --page raw --csv > gpurun_out/ncu_${TAG}_raw.csv 2>/dev/null
ncu -i /tmp/ncu_$TAG.ncu-rep --page source --csv > /tmp/ncu_${TAG}_source.csv 2>/dev/null
python scripts/ncu_summarize.py gpurun_out/ncu_${TAG}_raw.csv "$TAG" | head -20
python - <<PY
import csv
rows = list(csv.reader(open("/tmp/ncu_${TAG}_source.csv")))
hdr = next(i for i, r in enumerate(rows) if "Source" in r and any("Sampl" in c for c in r))
h = rows[hdr]
si = h.index("Source")
smp = next(i for i, c in enumerate(h) if c.startswith("# Samples") or c == "Warp Stall Sampling (All Samples)" or "Sampling (All" in c)
print("columns:", [c for c in h if "ampl" in c][:6])
body = [r for r in rows[hdr + 1:] if len(r) > smp]
tot = sum(float(r[smp] or 0) for r in body)
top = sorted(range(len(body)), key=lambda i: -float(body[i][smp] or 0))[:40]
print("total samples", tot)
for i in sorted(top):
    print(f"{i:6d} {float(body[i][smp] or 0) / tot:6.1%}  {body[i][si][:110]}")
PY
