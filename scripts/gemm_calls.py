#!/usr/bin/env python
"""Per-shape table of the gnnx_gemm calls of one training step (CUDA-event timed through _lib's timing hook)."""
import collections, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from gnnepcsaft_b200 import create_model, graph, ops, synth
from gnnepcsaft_b200._lib import lib
from gnnepcsaft_b200.train import Trainer

wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
dev = torch.device("cuda", 0)
ops.set_matmul_precision("tf32")
cfg = bench.WORKLOADS[wl]
target = "para" if cfg["num_para"] == 3 else "assoc"
graphs = bench.make_dataset(0, 4096)
store = graph.DeviceGraphStore(synth.pack(graphs), dev)
model = create_model(dict(cfg), store.calc_deg()).to(dev).train()
tr = Trainer(model, target=target)
ids = np.arange(4096)
for _ in range(2):
    tr.step(store.collate(ids))
lib().start_timing()
tr.step(store.collate(ids))
rec = lib().stop_timing()
tot = collections.defaultdict(lambda: [0, 0.0])
other = collections.defaultdict(lambda: [0, 0.0])
for name, a, ms in rec:
    if name == "gnnx_gemm":
        key = (a[8], a[9], a[10], a[2], a[5], a[12], bool(a[13]))
        tot[key][0] += 1; tot[key][1] += ms
    else:
        other[name][0] += 1; other[name][1] += ms
print(f"{'M':>7s} {'N':>5s} {'K':>6s} ta tb fl mk  n   ms/call  GB/s  TF/s")
for k, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    M, N, K, ta, tb, fl, mk = k
    byts = 4.0 * (M * K + N * K + M * N) + (4.0 * M * N if fl & 2 else 0) + (4.0 * M * N if mk else 0)
    print(f"{M:7d} {N:5d} {K:6d} {ta:2d} {tb:2d} {fl:2d} {int(mk):2d} {n:3d} {ms / n:8.3f} {byts / (ms / n) / 1e6:6.0f} {2.0 * M * N * K / (ms / n) / 1e9:6.0f}")
print("gemm total ms", sum(v[1] for v in tot.values()))
for k, (n, ms) in sorted(other.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:32s} x{n:3d} {ms:8.3f} ms")
