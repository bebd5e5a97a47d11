#!/usr/bin/env python
"""Diagnostic: where the degree-grouped PNA path (training mode, fp32 tier, cfg1) leaves the general path -- per-layer
conv and BatchNorm outputs of the two, in node order."""
import copy, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import models, ops, pna_fn, nn as gnn, graph
from tests.util import PNA_CFG1, deg_hist, make_batch

torch.set_float32_matmul_precision("highest"); ops.set_matmul_precision("fp32"); torch.manual_seed(0)
graphs, b = make_batch(32, seed=1234)
cfg = dict(PNA_CFG1, deg=deg_hist(graphs))
m1 = G.GNNePCSAFT(copy.deepcopy(cfg)).cuda().train()
args = [b[k].cuda() for k in ("x", "edge_index", "edge_attr", "batch")]
Gn = int(b["batch"][-1]) + 1

def run(grouped):
    m = copy.deepcopy(m1); pna_fn.GROUPED_POST = grouped
    taps = []
    hooks = []
    for l, (c, bn) in enumerate(zip(m.convs, m.batch_norms)):
        hooks.append(c.register_forward_hook(lambda mod, i, o, l=l: taps.append((f"conv{l}", o.detach().clone()))))
        hooks.append(bn.register_forward_hook(lambda mod, i, o, l=l: taps.append((f"bn{l}", o.detach().clone()))))
    try:
        out = m(*args, num_graphs=Gn).detach()
    finally:
        pna_fn.GROUPED_POST = True
        for h in hooks: h.remove()
    return out, taps, m

og, tg, mg = run(True)
o0, t0, m0 = run(False)
bk = graph.graph_index_for(args[1], args[0].shape[0]).degree_buckets()
print("buckets:", bk.bucket_ptr.tolist(), "capacity", bk.capacity, "N", args[0].shape[0])
for (ng, a), (n0, c) in zip(tg, t0):
    if a.shape != c.shape:                         # bucket order -> node order
        a = ops.gather_rows(a.contiguous(), bk.node_pos)
    d = (a.double() - c.double()).abs()
    print(f"{ng:6s} shape {tuple(c.shape)} max|d| {float(d.max()):.3e} scale {float(c.abs().max()):.3g}  worst col {int(d.max(0).values.argmax())} worst row {int(d.max(1).values.argmax())}")
print("pred max|d|", float((og - o0).abs().max()))
for (k, v), (_, v0) in zip(mg.state_dict().items(), m0.state_dict().items()):
    if "running" in k:
        print(f"{k:40s} max|d| {float((v.double() - v0.double()).abs().max()):.3e}  scale {float(v0.abs().max()):.3g}")
