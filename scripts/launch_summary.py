#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name."""
import collections, csv, re, sys
path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
tot, cnt = collections.defaultdict(float), collections.Counter()
for row in csv.DictReader(lines):
    name = re.sub(r"\(.*", "", row["Kernel Name"])
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    v = v / 1e3 if unit.startswith("n") else (v * 1e3 if unit.startswith("m") else v)
    tot[name] += v; cnt[name] += 1
T = sum(tot.values())
print(f"total {T:.1f} us over {sum(cnt.values())} launches")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:top]:
    print(f"{v:10.1f} us {100 * v / T:5.1f}%  x{cnt[k]:4d}  {v / cnt[k]:8.1f} us/launch  {k[:80]}")
