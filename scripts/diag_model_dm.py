import sys, os, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnnepcsaft_b200 import graph, ops
from oracle import pyg_restated as O
from tests.util import *
torch.set_float32_matmul_precision("highest")
graphs, b = make_batch(32, seed=1234)
cfg = dict(PNA_CFG1, deg=deg_hist(graphs))
cap = {}
def run(dtype):
    torch.manual_seed(1)
    mo = O.GNNePCSAFT(cfg).to(dtype)
    mod = mo.convs[2].aggr_module
    def hooked(x, index, dim_size):
        x.retain_grad()
        inner = torch.cat([O._AGGRS[a](x, index, dim_size) for a in mod.aggr_names], dim=-1)
        inner.retain_grad()
        deg = O.degree(index, dim_size, dtype=inner.dtype).view(-1, 1, 1)
        outs = [inner, inner * (torch.log(deg + 1) / mod.avg_deg_log), inner * (mod.avg_deg_log / torch.log(deg.clamp(min=1) + 1))]
        cap[dtype] = (x, inner, index)
        return torch.cat(outs, -1)
    mod.forward = hooked
    po = mo(b["x"], b["edge_index"], b["edge_attr"], b["batch"])
    lo, _ = O.training_loss(po, b['para'].to(dtype)); lo.backward()
run(torch.float32); run(torch.float64)
m32, inner32, idx = cap[torch.float32]; m64, inner64, _ = cap[torch.float64]
N = b["x"].shape[0]; E = idx.numel()
g = graph.GraphIndex.build(b["edge_index"].cuda(), N)
perm = g.csr_eid.cpu().long()
ms = m32.detach()[:, 0, :][perm].contiguous().cuda()
agg = ops.pna_aggregate_fwd(ms, g.rowptr, 1)
print("agg fwd max abs diff vs oracle32 inner:", (agg.cpu() - inner32.detach()[:, 0, :]).abs().max().item())
dagg = inner32.grad[:, 0, :].contiguous().cuda()
dm = ops.pna_aggregate_bwd(dagg, ms, agg, g.rowptr, 1)
dmo = torch.empty(E, 64); dmo[perm] = dm.cpu()
r32, r64 = m32.grad[:, 0, :], m64.grad[:, 0, :]
eg, er = (dmo.double() - r64).abs(), (r32.double() - r64).abs()
print(f"dm: gpu-vs-64 {eg.max():.3e}  ref32-vs-64 {er.max():.3e}  gpu-vs-32 {(dmo - r32).abs().max():.3e} scale {r64.abs().max():.3e}")
for sel, name in ((slice(0, 64), "mean"), (slice(64, 128), "min"), (slice(128, 192), "max"), (slice(192, 256), "std")):
    dz = torch.zeros_like(dagg); dz[:, sel] = dagg[:, sel]
    d1 = ops.pna_aggregate_bwd(dz, ms, agg, g.rowptr, 1)
    d1o = torch.empty(E, 64); d1o[perm] = d1.cpu()
    # oracle term: autograd of inner32 wrt m32 with the masked cotangent
    ct = torch.zeros_like(inner32); ct[:, 0, sel] = inner32.grad[:, 0, sel]
    (gref,) = torch.autograd.grad(inner32, m32, ct, retain_graph=True)
    i = int((d1o - gref[:, 0, :]).abs().argmax()); e, c = divmod(i, 64)
    d = int(idx[e]); vals = m32.detach()[idx == d][:, 0, c]
    print(f"  {name}: max diff {(d1o - gref[:, 0, :]).abs().max():.3e} (scale {gref.abs().max():.3e}) worst edge {e} col {c} cnt {vals.numel()} vals {vals.tolist()} gpu {d1o[e, c]:.6e} ref {gref[e, 0, c]:.6e}")
