#!/usr/bin/env python
"""Per-kernel summary of an `ncu -i X.ncu-rep --page raw --csv` dump: the counters the judge asks for (duration, DRAM
bytes and throughput, L2 hit rate, issue-active, achieved occupancy, tensor-pipe activity), averaged over the launches
of each kernel name.

    python scripts/ncu_summarize.py gpurun_out/ncu_r2a_kernels_raw.csv [title] > profiles/ncu_r2a_kernels_set_full.txt
"""
import collections, csv, re, sys

WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__registers_per_thread", "regs/thread"),
    ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("lts__t_bytes.sum", "L2 bytes"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm throughput %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "stall long scoreboard"),
    ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long scoreboard / issue"),
]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3,
         "msecond": 1e3}


def main():
    path = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else path
    stalls = "--stalls" in sys.argv
    with open(path) as f:
        rows = list(csv.reader(l for l in f if l.startswith('"')))
    head, units, body = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(head)}
    kname = col.get("Kernel Name")
    acc = collections.OrderedDict()
    for r in body:
        name = re.sub(r"\(.*", "", r[kname]).replace("void ", "").replace("gnnx::", "")
        a = acc.setdefault(name, collections.defaultdict(list))
        for m, _ in WANT:
            if m in col and r[col[m]] not in ("", "n/a"):
                try:
                    v = float(r[col[m]].replace(",", ""))
                except ValueError:
                    continue
                a[m].append(v * SCALE.get(units[col[m]], 1.0))
    print(title)
    for name, a in acc.items():
        n = len(a["gpu__time_duration.sum"])
        print(f"---- {name}  (x{n} launches, averages)")
        for m, label in WANT:
            if a[m]:
                v = sum(a[m]) / len(a[m])
                unit = "us" if m.startswith("gpu__time") else ("MB" if "bytes" in m else "")
                if "bytes" in m:
                    v /= 1e6
                print(f"   {label:34s} {v:14.3f} {unit}")
        if stalls:                                # every stall-reason / pipe-utilisation column of the first launch of this kernel
            first = next(r for r in body if re.sub(r"\(.*", "", r[kname]).replace("void ", "").replace("gnnx::", "") == name)
            rows_ = []
            for cname, i in col.items():
                if ("issue_stalled" in cname and cname.endswith("_per_warp_active.pct")) or "pipe_" in cname and cname.endswith("pct_of_peak_sustained_active") \
                        or cname.startswith("l1tex__t_sector_hit") or cname.startswith("l1tex__data_bank_conflicts") or "local" in cname and cname.endswith(".sum"):
                    try:
                        rows_.append((float(first[i].replace(",", "")), cname))
                    except ValueError:
                        pass
            for v, cname in sorted(rows_, reverse=True)[:28]:
                print(f"      {cname:90s} {v:12.3f}")
        t = sum(a["gpu__time_duration.sum"]) / max(n, 1)
        if a["dram__bytes_read.sum"] and t > 0:
            b = (sum(a["dram__bytes_read.sum"]) + sum(a["dram__bytes_write.sum"])) / n
            print(f"   {'dram GB/s (read+write / duration)':34s} {b / t / 1e3:14.1f}")


if __name__ == "__main__":
    main()
