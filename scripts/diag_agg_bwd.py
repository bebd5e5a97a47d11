import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gnnepcsaft_b200 import graph, ops
from oracle import pyg_restated as O
from tests.util import make_batch
torch.manual_seed(0)
T, F = 1, 64
_, b = make_batch(48, seed=9)
ei, N = b["edge_index"], b["x"].shape[0]
E, H = ei.shape[1], T * F
g = graph.GraphIndex.build(ei.cuda(), N)
perm = g.csr_eid.cpu().long()
m = torch.randn(E, H)
def oracle(dtype):
    mo = m.detach().clone().to(dtype).requires_grad_(True)
    outs = [O._AGGRS[a](mo.view(E, T, F), ei[1], N) for a in ("mean", "min", "max", "std")]
    return mo, outs
w = torch.randn(N, T, 4 * F)
res = {}
for name, sel in (("all", [0, 1, 2, 3]), ("mean", [0]), ("min", [1]), ("max", [2]), ("std", [3])):
    ww = torch.zeros_like(w)
    for s in sel:
        ww[..., s * F:(s + 1) * F] = w[..., s * F:(s + 1) * F]
    m32, o32 = oracle(torch.float32); m64, o64 = oracle(torch.float64)
    (torch.cat(o32, -1) * ww).sum().backward(); (torch.cat(o64, -1) * ww.double()).sum().backward()
    ms = m[perm].contiguous().cuda()
    agg = ops.pna_aggregate_fwd(ms, g.rowptr, T)
    dm = ops.pna_aggregate_bwd(ww.reshape(N, -1).contiguous().cuda(), ms, agg, g.rowptr, T)
    dmo = torch.empty_like(m); dmo[perm] = dm.cpu()
    eg = (dmo.double() - m64.grad).abs(); er = (m32.grad.double() - m64.grad).abs()
    print(f"{name:5s} gpu err max {eg.max():.3e}  ref32 err max {er.max():.3e}  scale {m64.grad.abs().max():.3e}")
    if name in ("all", "std"):
        i = int(eg.argmax()); e, c = divmod(i, H)
        d = int(ei[1, e]); cnt = int((ei[1] == d).sum())
        vals = m[ei[1] == d][:, c]
        print("   worst edge", e, "col", c, "dst", d, "cnt", cnt, "vals", vals.tolist(), "gpu", dmo[e, c].item(), "r32", m32.grad[e, c].item(), "r64", m64.grad[e, c].item())
        print("   agg(gpu) mean/min/max/std", [agg[d, s * F + c].item() for s in range(4)], "oracle32 std", o32[3][d, 0, c].item(), "oracle64 std", o64[3][d, 0, c].item())
