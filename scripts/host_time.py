#!/usr/bin/env python
"""Host issue time against device time of steady-state training steps: is the step GPU-bound or launch-bound?"""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from gnnepcsaft_b200 import create_model, graph, ops, synth
from gnnepcsaft_b200.train import Trainer
wl = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
dev = torch.device("cuda", 0)
ops.set_matmul_precision("tf32")
cfg = bench.WORKLOADS[wl]
graphs = bench.make_dataset(0, 8192)
store = graph.DeviceGraphStore(synth.pack(graphs), dev)
model = create_model(dict(cfg), store.calc_deg()).to(dev).train()
tr = Trainer(model, target="para" if cfg["num_para"] == 3 else "assoc")
rng = np.random.default_rng(0)
ids = [rng.permutation(8192)[:4096] for _ in range(4)]
for i in range(10):
    tr.step(store.collate(ids[i % 4]))
torch.cuda.synchronize()
K = 30
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for i in range(K):
    tr.step(store.collate(ids[i % 4]))
e1.record(); t_host = time.perf_counter() - t0
torch.cuda.synchronize(); t_all = time.perf_counter() - t0
# un-throttled host cost: right after a synchronisation the launch queue is empty, so one step's issue time is pure host work
hs = []
for i in range(6):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    tr.step(store.collate(ids[i % 4]))
    hs.append(time.perf_counter() - t0)
print(f"{wl}: host issue of ONE step into an empty queue: median {1e3 * sorted(hs)[len(hs) // 2]:.2f} ms (min {1e3 * min(hs):.2f})")
print(f"{wl}: host issue {1e3 * t_host / K:.2f} ms/step, device {e0.elapsed_time(e1) / K:.2f} ms/step, wall {1e3 * t_all / K:.2f} ms/step")
