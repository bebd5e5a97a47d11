"""Diagnostic: per-parameter gradient error of the CUDA path and of the fp32 oracle, both against the fp64 oracle."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import ops
from oracle import pyg_restated as O
from tests.util import *

torch.set_float32_matmul_precision("highest")
graphs, b = make_batch(32, seed=1234)
for cfg, target in [(PNA_CFG1, 'para'), (GAT_CFG_SMALL, 'assoc')]:
    cfg = dict(cfg, deg=deg_hist(graphs))
    torch.manual_seed(1)
    mo = O.GNNePCSAFT(cfg); mo64 = copy.deepcopy(mo).double()
    mp = G.GNNePCSAFT(copy.deepcopy(cfg)).cuda(); mp.load_state_dict(mo.state_dict())
    po = mo(b["x"], b["edge_index"], b["edge_attr"], b["batch"])
    p64 = mo64(b["x"], b["edge_index"], b["edge_attr"], b["batch"])
    pp = mp(b["x"].cuda(), b["edge_index"].cuda(), b["edge_attr"].cuda(), b["batch"].cuda())
    print("fwd err gpu %.3e ref32 %.3e" % ((pp.double().cpu()-p64).abs().max(), (po.double()-p64).abs().max()))
    lo, _ = O.training_loss(po, b[target]); l64, _ = O.training_loss(p64, b[target].double())
    lo.backward(); l64.backward()
    out2, dpred = ops.relhuber_fwd_bwd(pp.detach().contiguous(), b[target].cuda())
    pp.backward(dpred)
    d64, d32 = dict(mo64.named_parameters()), dict(mo.named_parameters())
    gmax = max(float(p.grad.abs().max()) for p in mo64.parameters())
    print("global grad scale %.3e" % gmax)
    for n, p in mp.named_parameters():
        g64 = d64[n].grad
        eg = float((p.grad.double().cpu() - g64).abs().max()); er = float((d32[n].grad.double() - g64).abs().max())
        flag = "  <<<" if eg > 4 * er and eg > 1e-6 * gmax else ""
        print(f"{n:45s} scale {float(g64.abs().max()):.2e} gpu {eg:.2e} ref32 {er:.2e}{flag}")
