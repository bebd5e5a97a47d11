#!/usr/bin/env python
"""Generates tests/golden/*.npz: seeded inputs, weights, and the CPU oracle's outputs for them.

The reference (torch_geometric / ogb) cannot be imported in this container, so these vectors come from the
restated oracle (oracle/pyg_restated.py) -- they pin the ORACLE against accidental change and give the GPU tests
a fixture that does not need the oracle at run time; they do not upgrade the parity status ("parity unpinned").

    python scripts/make_golden.py        # rewrites tests/golden/
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gnnepcsaft_b200 import synth  # noqa: E402  (host-side generator only; no CUDA needed)
from oracle import pyg_restated as O  # noqa: E402

CASES = {
    "pna_h32_l2": (dict(conv="PNA", hidden_dim=32, propagation_depth=2, dropout=0.0, global_pool="add", num_para=3, towers=2,
                        pre_layers=2, post_layers=2, heads=2, add_self_loops=True, model="gnn"), "para", 16, 4321),
    "gatv2_h32_l2": (dict(conv="GATv2", hidden_dim=32, propagation_depth=2, dropout=0.0, global_pool="add", num_para=2,
                          towers=1, pre_layers=2, post_layers=2, heads=2, add_self_loops=True, model="gnn"), "assoc", 16, 4322),
}


def main():
    torch.set_float32_matmul_precision("highest")
    torch.set_num_threads(1)
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, (cfg, target, num_graphs, seed) in CASES.items():
        graphs = synth.make_graphs(num_graphs, seed, 4, 18, 0.1)
        x, ei, ea, batch, para, assoc = synth.collate_numpy(graphs)
        t = torch.from_numpy
        og = [O.Graph(t(g.x), t(g.edge_index), t(g.edge_attr)) for g in graphs]
        deg = O.calc_deg(og)
        torch.manual_seed(seed)
        model = O.GNNePCSAFT(dict(cfg, deg=deg)).train()
        tgt = para if target == "para" else assoc
        pred = model(t(x), t(ei), t(ea), t(batch))
        loss, mape = O.training_loss(pred, t(tgt))
        state = {k: v.detach().clone().numpy() for k, v in model.state_dict().items()}     # BEFORE running stats matter: saved post-forward below
        loss.backward()
        csr = O.csr_reference(t(ei), x.shape[0])
        arrays = dict(x=x, edge_index=ei, edge_attr=ea, batch=batch, target=tgt, deg=np.asarray(deg, np.int64),
                      pred=pred.detach().numpy(), loss=np.float32(loss.item()), mape=np.float32(mape.item()),
                      rowptr=csr["rowptr"].numpy(), eid=csr["eid"].numpy(),
                      config=np.frombuffer(json.dumps(dict(cfg, target=target)).encode(), dtype=np.uint8))
        torch.manual_seed(seed)
        fresh = O.GNNePCSAFT(dict(cfg, deg=deg))                                           # pre-forward state (running stats untouched)
        for k, v in fresh.state_dict().items():
            arrays["state/" + k] = v.numpy()
        for k, p in model.named_parameters():
            arrays["grad/" + k] = p.grad.numpy()
        for k, v in state.items():
            if "running_" in k:
                arrays["after/" + k] = v
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **arrays)
        print(f"{name}: N={x.shape[0]} E={ei.shape[1]} loss={loss.item():.6f} -> {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
