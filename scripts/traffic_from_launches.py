#!/usr/bin/env python
"""DRAM traffic per entry-point call from an ncu launch list (gpu__time_duration + dram__bytes_{read,write}.sum):

    python scripts/traffic_from_launches.py profiles/launches_r1f_cfg2_tf32.csv cfg2 2 profiles/traffic_r1.json

`steps` = training steps inside the capture (profile_step.py runs 1 warm-up + N).  bench.py reads the JSON to fill
`roofline.traffic` (bytes per call of the dominant entry point, same unit as `roofline.bytes_per_launch`)."""
import collections, csv, json, os, re, sys

KERNEL_TO_ENTRY = [
    (r"tc::k_gemm_tf32|k_splitk_reduce|k_gemm_f32|k_pack_relu_bits|k_group_reduce", "gnnx_gemm"),
    (r"k_pna_aggregate_bwd", "gnnx_pna_aggregate_bwd"), (r"k_pna_aggregate_fwd", "gnnx_pna_aggregate_fwd"),
    (r"k_gather_rows", "gnnx_gather_rows"),
    (r"k_pna_gather_bwd|k_permute_rows", "gnnx_pna_gather_bwd"), (r"k_pna_gather_fwd", "gnnx_pna_gather_fwd"),
    (r"k_pna_post_combine_bwd", "gnnx_pna_post_combine_bwd"), (r"k_pna_post_combine_fwd", "gnnx_pna_post_combine_fwd"),
    (r"k_bn_bwd", "gnnx_bn_relu_bwd"), (r"k_bn_stats|k_bn_apply", "gnnx_bn_relu_fwd"),
    (r"k_gatv2_fwd", "gnnx_gatv2_fwd"), (r"k_gatv2_bwd|k_gatv2_datt", "gnnx_gatv2_bwd"),
    (r"k_tconv_fwd", "gnnx_tconv_fwd"), (r"k_tconv_bwd", "gnnx_tconv_bwd"),
    (r"k_code_sum|k_sum_partials", "gnnx_code_segment_sum"), (r"k_colsum", "gnnx_colsum"),
]
CALLS_PER_KERNEL = {"gnnx_gemm": None}     # gemm calls are counted by bench.py (several kernels per call)


def main():
    path, workload, steps, out = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    dram, launches = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        name = row["Kernel Name"]
        entry = next((e for pat, e in KERNEL_TO_ENTRY if re.search(pat, name)), None)
        if entry is None:
            continue
        v = float(row["Metric Value"].replace(",", ""))
        if row["Metric Name"].startswith("dram__bytes"):
            dram[entry] += v * scale.get(row["Metric Unit"], 1.0)
        elif row["Metric Name"].startswith("gpu__time_duration"):
            launches[entry] += 1
    doc = json.load(open(out)) if os.path.exists(out) else {}
    doc[workload] = {"source": os.path.relpath(path), "steps_in_capture": steps,
                     "dram_bytes_per_step": {e: dram[e] / steps for e in sorted(dram)},
                     "kernel_launches_per_step": {e: launches[e] / steps for e in sorted(launches)}}
    json.dump(doc, open(out, "w"), indent=1)
    for e in sorted(dram, key=lambda k: -dram[k]):
        print(f"{e:28s} {dram[e] / steps / 1e6:10.1f} MB/step  {launches[e] / steps:6.1f} kernel launches/step")


if __name__ == "__main__":
    main()
