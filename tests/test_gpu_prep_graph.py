"""The parameter-only work of all PNAConv layers replayed from ONE CUDA graph (GNNePCSAFT._prepare_convs_graph) against
the same work done inline by every layer: training several steps on changing batches must give bit-identical losses,
parameters and optimizer state; the graph must really have been captured and replayed; and a change of what the graph
reads through fixed addresses (parameter storage replaced, another degree-group count) must re-key it, not reuse it."""
import copy

import numpy as np
import pytest
import torch

import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import graph, ops, streams, synth
from gnnepcsaft_b200.train import Trainer
from tests.util import PNA_CFG1, deg_hist, make_batch

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def _restore():
    was = (streams.PREP_AHEAD, streams.PREP_GRAPH)
    yield
    streams.PREP_AHEAD, streams.PREP_GRAPH = was
    ops.set_matmul_precision("fp32")


def _setup(prec):
    ops.set_matmul_precision(prec)
    graphs, _ = make_batch(384, seed=91)
    cfg = dict(PNA_CFG1, hidden_dim=128, deg=deg_hist(graphs))
    store = graph.DeviceGraphStore(synth.pack(graphs), torch.device(DEV))
    torch.manual_seed(5)
    model = G.GNNePCSAFT(copy.deepcopy(cfg)).to(DEV).train()
    return model, Trainer(model), store, cfg


def _run(tr, store, steps, rng):
    out = []
    for _ in range(steps):
        ids = rng.permutation(store.num_graphs)[:256]
        out.append(tr.step(store.collate(ids)).clone())
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_graph_replayed_preparation_equals_inline_preparation(prec):
    streams.PREP_AHEAD, streams.PREP_GRAPH = True, True
    m1, t1, store, _ = _setup(prec)
    l1 = _run(t1, store, 6, np.random.default_rng(0))
    ent = m1._prep_graph
    assert ent is not None and ent["graph"] is not None and ent["launches"] > 0, "the graph must have been captured (step 2) and replayed"
    streams.PREP_AHEAD = False
    m2, t2, store2, _ = _setup(prec)
    l2 = _run(t2, store2, 6, np.random.default_rng(0))
    assert m2._prep_graph is None
    for i, (a, b) in enumerate(zip(l1, l2)):
        assert torch.equal(a, b), f"loss of step {i}: {a.tolist()} vs {b.tolist()}"
    assert torch.equal(t1.flat.data, t2.flat.data) and torch.equal(t1.exp_avg, t2.exp_avg) and torch.equal(t1.max_exp_avg_sq, t2.max_exp_avg_sq)


def test_graph_is_rekeyed_when_its_fixed_addresses_or_shapes_change():
    streams.PREP_AHEAD, streams.PREP_GRAPH = True, True
    model, tr, store, cfg = _setup("fp32")
    _run(tr, store, 3, np.random.default_rng(1))
    first = model._prep_graph
    assert first["graph"] is not None
    # a parameter moves to new storage (what load-with-assign or a dtype / device round trip does): the captured graph
    # still reads the old address and must not be replayed
    p = next(model.convs[0].parameters())
    with torch.no_grad():
        p.data = p.data.clone()
    ref = G.GNNePCSAFT(copy.deepcopy(cfg)).to(DEV).train()      # same values, its own storage, inline preparation below
    ref.load_state_dict(model.state_dict())
    bt = store.collate(np.arange(256))
    out = model.forward_batch(bt)
    assert model._prep_graph is not first, "new parameter storage must re-key the graph"
    streams.PREP_AHEAD = False
    want = ref.forward_batch(store.collate(np.arange(256)))
    assert torch.equal(out, want)
    # another degree-group count (no dataset hint: the batch's own maximum in-degree decides) re-keys as well
    streams.PREP_AHEAD = True
    model.degree_hint = None
    small = store.collate(np.arange(8))
    big = store.collate(np.arange(256))
    for b in (small, big, small, big):
        model.forward_batch(b)
    torch.cuda.synchronize()
    streams.PREP_AHEAD = False
    ref.degree_hint = None
    assert torch.equal(model.forward_batch(store.collate(np.arange(8))), ref.forward_batch(store.collate(np.arange(8))))
