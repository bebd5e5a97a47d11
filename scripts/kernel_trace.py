#!/usr/bin/env python
"""In-situ per-kernel device timeline of training steps (lib().trace_begin / trace_end: CUDA events around every gnnx
kernel launch; warm caches and real predecessors, unlike an ncu launch list, which serialises and flushes).

    python scripts/kernel_trace.py [--workload cfg2] [--steps 2] [--out gpurun_out/trace.txt]
"""
import argparse, collections, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from gnnepcsaft_b200 import create_model, graph, ops, synth
from gnnepcsaft_b200._lib import lib
from gnnepcsaft_b200.train import Trainer

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="cfg2")
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--out", default=None)
args = ap.parse_args()
dev = torch.device("cuda", 0)
ops.set_matmul_precision("tf32")
cfg = bench.WORKLOADS[args.workload]
target = "para" if cfg["num_para"] == 3 else "assoc"
graphs = bench.make_dataset(0, 4096)
store = graph.DeviceGraphStore(synth.pack(graphs), dev)
model = create_model(dict(cfg), store.calc_deg()).to(dev).train()
tr = Trainer(model, target=target)
ids = np.arange(4096)
for _ in range(3):
    tr.step(store.collate(ids))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(args.steps):
    tr.step(store.collate(ids))
e1.record()
torch.cuda.synchronize()
plain_ms = e0.elapsed_time(e1) / args.steps
lib().trace_begin()
e0.record()
for _ in range(args.steps):
    tr.step(store.collate(ids))
e1.record()
rows = lib().trace_end()
traced_ms = e0.elapsed_time(e1) / args.steps
agg = collections.OrderedDict()
for name, us, gap in rows:
    short = name.split("(")[0].replace("void ", "").replace("gnnx::", "")
    a = agg.setdefault(short, [0, 0.0, 0.0])
    a[0] += 1; a[1] += us; a[2] += gap
tot_k = sum(a[1] for a in agg.values()); tot_g = sum(a[2] for a in agg.values())
lines = [f"{args.workload}: {args.steps} steps, {len(rows)} gnnx launches; step {plain_ms:.3f} ms untraced, {traced_ms:.3f} ms traced; "
         f"per step: gnnx kernels {tot_k / args.steps / 1e3:.3f} ms, between gnnx kernels (other kernels + gaps) {tot_g / args.steps / 1e3:.3f} ms",
         f"{'kernel':44s} {'n/step':>7s} {'us/launch':>10s} {'ms/step':>8s} {'share':>6s} {'gap before us':>14s}"]
for k, (n, us, gap) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    lines.append(f"{k[:44]:44s} {n / args.steps:7.1f} {us / n:10.2f} {us / args.steps / 1e3:8.3f} {us / tot_k:6.1%} {gap / n:14.2f}")
txt = "\n".join(lines)
print(txt)
if args.out:
    open(args.out, "w").write(txt + "\n")
