#!/usr/bin/env python
"""Diagnostic: the custom-op face (general scaler-folded PNA path, recomputation backward) against the fused face with
and without the degree-grouped path, and both against the fp64 oracle, in training mode at cfg1."""
import copy, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import models, ops, pna_fn
from oracle import pyg_restated as O
from tests.util import PNA_CFG1, deg_hist, make_batch

torch.set_float32_matmul_precision("highest"); ops.set_matmul_precision("fp32"); torch.manual_seed(0)
graphs, b = make_batch(32, seed=1234)
cfg = dict(PNA_CFG1, deg=deg_hist(graphs))
m1 = G.GNNePCSAFT(copy.deepcopy(cfg)).cuda().train()
mo = O.GNNePCSAFT(copy.deepcopy(cfg)).double().train(); mo.load_state_dict({k: v.double() if v.is_floating_point() else v for k, v in m1.state_dict().items()})
args = [b[k].cuda() for k in ("x", "edge_index", "edge_attr", "batch")]
Gn = int(b["batch"][-1]) + 1
ref = mo(b["x"], b["edge_index"], b["edge_attr"], b["batch"]).detach()
def run(face, grouped):
    m = copy.deepcopy(m1); pna_fn.GROUPED_POST = grouped; models.use_custom_ops(face)
    try: return m(*args, num_graphs=Gn).detach().double().cpu()
    finally: models.use_custom_ops(False); pna_fn.GROUPED_POST = True
outs = {"fused grouped": run(False, True), "fused general": run(False, False), "custom-op": run(True, True), "fused grouped again": run(False, True)}
for k, v in outs.items():
    print(f"{k:22s} max|d| vs fp64 oracle {float((v - ref).abs().max()):.3e}  (scale {float(ref.abs().max()):.3g})")
ks = list(outs)
for i in range(len(ks)):
    for j in range(i + 1, len(ks)):
        print(f"{ks[i]:22s} vs {ks[j]:22s} max|d| {float((outs[ks[i]] - outs[ks[j]]).abs().max()):.3e}")
