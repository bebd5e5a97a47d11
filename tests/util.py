"""Shared helpers for the parity tests: synthetic batches as oracle (CPU) and product (CUDA) inputs."""
import numpy as np
import torch

from gnnepcsaft_b200 import synth
from oracle import pyg_restated as O

PNA_CFG1 = dict(conv="PNA", hidden_dim=64, propagation_depth=3, dropout=0.0, global_pool="add", num_para=3,
                towers=1, pre_layers=2, post_layers=4, heads=2, add_self_loops=True, model="gnn")
GAT_CFG_SMALL = dict(conv="GATv2", hidden_dim=64, propagation_depth=3, dropout=0.0, global_pool="add", num_para=2,
                     towers=1, pre_layers=2, post_layers=4, heads=2, add_self_loops=True, model="gnn")


def make_batch(num_graphs=32, seed=1234, n_min=10, n_max=40, single_atom_frac=0.05):
    graphs = synth.make_graphs(num_graphs, seed, n_min, n_max, single_atom_frac)
    x, ei, ea, batch, para, assoc = synth.collate_numpy(graphs)
    t = torch.from_numpy
    return graphs, dict(x=t(x), edge_index=t(ei), edge_attr=t(ea), batch=t(batch), para=t(para), assoc=t(assoc))


def oracle_graphs(graphs):
    t = torch.from_numpy
    return [O.Graph(t(g.x), t(g.edge_index), t(g.edge_attr), para=t(g.para), assoc=t(g.assoc)) for g in graphs]


def deg_hist(graphs):
    return O.calc_deg(oracle_graphs(graphs))


def rel_err(a: torch.Tensor, b: torch.Tensor, atol=1e-6):
    """max |a-b| / (|b| + atol-scale): relative error with an absolute floor for near-zero entries."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    scale = b.abs().max().clamp(min=1e-30)
    return float(((a - b).abs() / (b.abs() + atol + 1e-4 * scale)).max()) if a.numel() else 0.0


def assert_close(a, b, rtol=1e-5, atol=1e-6, what=""):
    """|a-b| <= atol + rtol*|b| elementwise, with |b| floored at 1e-2 of the tensor's scale (entries that are
    sums of O(scale) terms carry O(eps*scale) rounding regardless of their own magnitude)."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    if a.numel() == 0:
        return
    scale = float(b.abs().max())
    tol = atol + rtol * torch.maximum(b.abs(), torch.full_like(b, 1e-2 * scale))
    bad = (a - b).abs() > tol
    if bad.any():
        i = int(((a - b).abs() / tol).argmax())
        raise AssertionError(f"{what}: {int(bad.sum())}/{a.numel()} entries out of tolerance; worst at flat index {i}: "
                             f"got {a.flatten()[i]:.9g} want {b.flatten()[i]:.9g} (scale {scale:.3g})")
