"""Layer-by-layer comparison of the CUDA model against the oracle on a golden fixture (diagnostic)."""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import ops, functional as Fn
from oracle import pyg_restated as O

name = sys.argv[1] if len(sys.argv) > 1 else "pna_h32_l2"
z = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
cfg = json.loads(bytes(z["config"]).decode()); cfg.pop("target")
torch.set_float32_matmul_precision("highest"); ops.set_matmul_precision("fp32")
t = torch.from_numpy
state = {k[6:]: t(z[k]) for k in z.files if k.startswith("state/")}
mo = O.GNNePCSAFT(dict(cfg, deg=z["deg"].tolist())).train(); mo.load_state_dict(state)
mp = G.create_model(dict(cfg), z["deg"].tolist()).cuda().train(); mp.load_state_dict(state)
acts_o, acts_p = {}, {}
def hook(store, n):
    def f(mod, inp, out):
        store[n] = out.detach().cpu() if isinstance(out, torch.Tensor) else None
    return f
for (n, m) in mo.named_modules():
    if n and n.count(".") <= 1: m.register_forward_hook(hook(acts_o, n))
for (n, m) in mp.named_modules():
    if n and n.count(".") <= 1: m.register_forward_hook(hook(acts_p, n))
po = mo(t(z["x"]), t(z["edge_index"]), t(z["edge_attr"]), t(z["batch"]))
pp = mp(t(z["x"]).cuda(), t(z["edge_index"]).cuda(), t(z["edge_attr"]).cuda(), t(z["batch"]).cuda())
print("oracle vs golden", float((po.detach() - t(z["pred"])).abs().max()))
print("cuda   vs golden", float((pp.detach().cpu() - t(z["pred"])).abs().max()))
for n in acts_o:
    a, b = acts_o.get(n), acts_p.get(n)
    if a is None or b is None or a.shape != b.shape:
        print(f"{n:30s} oracle {None if a is None else tuple(a.shape)} cuda {None if b is None else tuple(b.shape)}"); continue
    print(f"{n:30s} {tuple(a.shape)} max|diff| {float((a - b).abs().max()):.3e} scale {float(a.abs().max()):.3e}")
# head pieces by hand
x = acts_o.get("global_pool")
if x is not None:
    m = mp.mlp
    xp = x.cuda()
    y0 = Fn.linear(xp, m[0].weight, m[0].bias); r0 = torch.nn.functional.linear(x, mo.mlp[0].weight, mo.mlp[0].bias)
    print("head lin0", float((y0.cpu() - r0).abs().max()))
