#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv` launch
list per kernel name: total time, share, launches, time per launch and (when captured) DRAM bytes per launch."""
import collections, csv, re, sys
path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
tot, cnt = collections.defaultdict(float), collections.Counter()
dram = collections.defaultdict(float)
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for row in csv.DictReader(lines):
    name = re.sub(r"\(.*", "", row["Kernel Name"])
    v = float(row["Metric Value"].replace(",", ""))
    unit, metric = row["Metric Unit"], row["Metric Name"]
    if metric.startswith("gpu__time_duration"):
        v = v / 1e3 if unit.startswith("n") else (v * 1e3 if unit.startswith("m") else v)
        tot[name] += v; cnt[name] += 1
    elif metric.startswith("dram__bytes"):
        dram[name] += v * scale.get(unit, 1.0)
T = sum(tot.values())
print(f"total {T:.1f} us over {sum(cnt.values())} launches")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:top]:
    extra = f"  dram {dram[k] / cnt[k] / 1e6:8.2f} MB/launch ({dram[k] / v / 1e3:6.0f} GB/s)" if dram else ""
    print(f"{v:10.1f} us {100 * v / T:5.1f}%  x{cnt[k]:4d}  {v / cnt[k]:8.1f} us/launch{extra}  {k[:70]}")
