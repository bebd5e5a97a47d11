"""Bulk screening (BASELINE config 5): node-balanced batches, shards with no communication, eval-mode parity.

Tolerance: fp32 tier, rel 1e-4 / atol 1e-5 against the CPU oracle's ``pred_with_bounds`` (eval mode: every graph is
independent of its batch, so any batching must give the oracle's per-molecule answer)."""
import copy

import numpy as np
import pytest
import torch

import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import graph, ops, synth
from gnnepcsaft_b200.infer import BulkScreener
from oracle import pyg_restated as O
from tests.util import PNA_CFG1, assert_close, deg_hist

pytestmark = pytest.mark.gpu


def test_sharded_screening_matches_oracle_and_unsharded():
    ops.set_matmul_precision("fp32")
    torch.set_float32_matmul_precision("highest")
    graphs = synth.make_inference_graphs(300, seed=11)
    cfg = dict(PNA_CFG1, deg=deg_hist(graphs))
    torch.manual_seed(3)
    mo = O.GNNePCSAFT(cfg).eval()
    with torch.no_grad():                              # non-trivial running statistics
        for bn in list(mo.batch_norms) + [mo.mlp[1], mo.mlp[4]]:
            m = bn.module if hasattr(bn, "module") else bn
            m.running_mean.normal_(0, 0.1); m.running_var.uniform_(0.5, 1.5)
    mp = G.GNNePCSAFT(copy.deepcopy(cfg)).cuda().eval()
    mp.load_state_dict(mo.state_dict(), strict=True)
    t = torch.from_numpy
    with torch.no_grad():
        ref = mo.pred_with_bounds(O.collate([O.Graph(t(g.x), t(g.edge_index), t(g.edge_attr)) for g in graphs]))
    store = graph.DeviceGraphStore(synth.pack(graphs), torch.device("cuda", 0))
    whole = BulkScreener(mp, store, max_nodes=2000, max_graphs=64)
    assert len(whole.batches) > 5 and whole.batches[0][0] == 0 and whole.batches[-1][1] == 300
    out = whole.run()
    assert_close(out, ref, 1e-4, 1e-5, "screening vs oracle")
    lo_b, hi_b = torch.tensor([1.0, 1.9, 50.0]), torch.tensor([25.0, 4.5, 550.0])
    assert bool((out.cpu() >= lo_b).all()) and bool((out.cpu() <= hi_b).all())
    parts = [BulkScreener(mp, store, r, 3, max_nodes=1500, max_graphs=50).run() for r in range(3)]
    assert sum(p.shape[0] for p in parts) == 300
    assert_close(torch.cat(parts), ref, 1e-4, 1e-5, "3 shards vs oracle")
    assert not mp.training
    # the streamed variant (pinned host array filled on a side stream, one wait at the end) gives the same numbers
    from gnnepcsaft_b200.infer import screen_to_host
    host, nbytes = screen_to_host(mp, store, max_nodes=2000, max_graphs=64)
    assert host.is_pinned() and nbytes == 300 * 3 * 4
    assert torch.equal(host, out.cpu())
    h1, _ = screen_to_host(mp, store, 1, 3, max_nodes=1500, max_graphs=50)
    assert torch.equal(h1, parts[1].cpu())


def test_output_stage_assembles_the_feos_parameter_vector(tmp_path):
    """both models' clipped predictions -> [m, sigma, eps/k, kappa_AB, eps_AB/k, mu, na, nb, MW],
    checked against the reference's formula (demo/utils.py:664-691) evaluated on the oracle's predictions."""
    from gnnepcsaft_b200.infer import screen_parameters
    from tests.util import GAT_CFG_SMALL
    ops.set_matmul_precision("fp32")
    torch.set_float32_matmul_precision("highest")
    graphs = synth.make_inference_graphs(120, seed=12)
    deg = deg_hist(graphs)
    torch.manual_seed(5)
    mo_p = O.GNNePCSAFT(dict(PNA_CFG1, deg=deg)).eval()
    mo_a = O.GNNePCSAFT(dict(GAT_CFG_SMALL, deg=deg)).eval()
    mp_p = G.GNNePCSAFT(dict(PNA_CFG1, deg=deg)).cuda().eval()
    mp_a = G.GNNePCSAFT(dict(GAT_CFG_SMALL, deg=deg)).cuda().eval()
    mp_p.load_state_dict(mo_p.state_dict(), strict=True)
    mp_a.load_state_dict(mo_a.state_dict(), strict=True)
    t = torch.from_numpy
    rng = np.random.default_rng(0)
    munanb = t(np.stack([rng.uniform(0, 3, 120), rng.integers(0, 3, 120), rng.integers(0, 3, 120)], 1).astype(np.float32))
    mw = t(rng.uniform(16, 400, 120).astype(np.float32))
    with torch.no_grad():
        b = O.collate([O.Graph(t(g.x), t(g.edge_index), t(g.edge_attr)) for g in graphs])
        msigmae, la = mo_p.pred_with_bounds(b), mo_a.pred_with_bounds(b)
    want = torch.cat([msigmae, 10 ** (la * torch.tensor([-1.0, 1.0])), munanb, mw[:, None]], 1)
    store = graph.DeviceGraphStore(synth.pack(graphs), torch.device("cuda", 0))
    got = screen_parameters(mp_p, store, munanb, mw, model_assoc=mp_a, max_nodes=1500, max_graphs=40,
                            out_path=str(tmp_path / "params"))
    assert got.shape == (120, 9)
    assert_close(got, want, 1e-4, 1e-5, "9-vector", floor=0.0)
    assert np.array_equal(np.load(str(tmp_path / "params") + ".rank0.npy"), got.cpu().numpy())
    # non-associating path: known association values (or zeros) pass straight through; shards concatenate
    known = t(rng.uniform(0, 1, (120, 2)).astype(np.float32))
    parts = [screen_parameters(mp_p, store, munanb, mw, assoc=known, rank=r, world=2) for r in range(2)]
    want2 = torch.cat([msigmae, known, munanb, mw[:, None]], 1)
    assert_close(torch.cat(parts), want2, 1e-4, 1e-5, "9-vector, known assoc", floor=0.0)
    zeros = screen_parameters(mp_p, store, munanb, mw)
    assert bool((zeros[:, 3:5] == 0).all())
    with pytest.raises(ValueError):
        screen_parameters(mp_p, store, munanb, mw, model_assoc=mp_a, assoc=known)
