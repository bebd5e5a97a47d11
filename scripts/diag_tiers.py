"""Debug aid: per-parameter gradient agreement of the tcgen05 tf32 tier with the CUDA-core fp32 tier (and, at sizes
the CPU finishes in seconds, with the fp64 oracle and with a bf16-product oracle = the reference's "medium" setting).

    python scripts/diag_tiers.py [graphs=4096] [hidden=256] [layers=6] [oracle=0|1]
"""
import copy
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gnnepcsaft_b200 as G                                     # noqa: E402
from gnnepcsaft_b200 import graph, ops, synth                    # noqa: E402

ng = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
H = int(sys.argv[2]) if len(sys.argv) > 2 else 256
L = int(sys.argv[3]) if len(sys.argv) > 3 else 6
with_oracle = len(sys.argv) > 4 and sys.argv[4] == "1"
cfg = dict(conv="PNA", hidden_dim=H, propagation_depth=L, dropout=0.0, global_pool="add", num_para=3,
           towers=1, pre_layers=2, post_layers=4, heads=2, add_self_loops=True, model="gnn")
dev = torch.device("cuda", 0)
graphs = synth.make_graphs(ng, seed=2235)
store = graph.DeviceGraphStore(synth.pack(graphs), dev)
deg = store.calc_deg()
b = store.collate(np.arange(ng))


def run(prec):
    ops.set_matmul_precision(prec)
    torch.manual_seed(0)
    m = G.create_model(copy.deepcopy(cfg), deg).to(dev).train()
    y = m.forward_batch(b)
    out, dy = ops.relhuber_fwd_bwd(y.detach().contiguous(), b.para)
    y.backward(dy)
    torch.cuda.synchronize()
    return m, y.detach(), {n: p.grad.detach().clone() for n, p in m.named_parameters()}


m32, y32, g32 = run("fp32")
mtf, ytf, gtf = run("tf32")
refs = {}
if with_oracle:
    from oracle import pyg_restated as O
    t = torch.from_numpy
    og = [O.Graph(t(g.x), t(g.edge_index), t(g.edge_attr), para=t(g.para), assoc=t(g.assoc)) for g in graphs]
    ob = O.collate(og)
    for name, dt in (("fp64", torch.float64), ("fp32cpu", torch.float32)):
        mo = O.GNNePCSAFT(dict(cfg, deg=deg)).train()
        mo.load_state_dict({k: v.cpu() for k, v in m32.state_dict().items()})
        mo = mo.to(dt)
        p = mo(ob.x, ob.edge_index, ob.edge_attr, ob.batch)
        loss, _ = O.training_loss(p, ob.para.to(dt))
        loss.backward()
        refs[name] = (p.detach(), {n: q.grad.detach() for n, q in mo.named_parameters()})
    y64 = refs["fp64"][0]
    for nm, y in (("cuda fp32", y32), ("cuda tf32", ytf), ("cpu fp32", refs["fp32cpu"][0])):
        print(f"pred {nm:10s} vs fp64 oracle: {float((y.cpu().double() - y64).norm() / y64.norm()):.3e}")

print(f"pred tf32 vs fp32 tier: {float((ytf - y32).norm() / y32.norm()):.3e}")
tot_d = tot_n = 0.0
rows = []
for n in g32:
    d, s = float((gtf[n] - g32[n]).norm()), float(g32[n].norm())
    tot_d += d * d; tot_n += s * s
    extra = ""
    if with_oracle:
        r = refs["fp64"][1][n]
        extra = (f"  |fp32-fp64| {float((g32[n].cpu().double() - r).norm() / max(float(r.norm()), 1e-30)):.2e}"
                 f"  |tf32-fp64| {float((gtf[n].cpu().double() - r).norm() / max(float(r.norm()), 1e-30)):.2e}"
                 f"  |cpu32-fp64| {float((refs['fp32cpu'][1][n].double() - r).norm() / max(float(r.norm()), 1e-30)):.2e}")
    rows.append((d, f"{n:45s} |g| {s:10.3e}  |tf32-fp32| {d:10.3e}  rel {d / max(s, 1e-30):9.2e}{extra}"))
for _, line in sorted(rows, key=lambda r: -r[0])[:40]:
    print(line)
print(f"all gradients: rel {np.sqrt(tot_d / tot_n):.3e}")
