"""GPU collator parity (integer work: BIT-EXACT against the oracle)."""
import numpy as np
import pytest
import torch

from gnnepcsaft_b200 import graph, ops, synth
from oracle import pyg_restated as O
from tests.util import make_batch, oracle_graphs

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _check_csr(ei: torch.Tensor, n: int):
    g = graph.GraphIndex.build(ei.to(DEV), n, check=True)
    torch.cuda.synchronize()
    ref = O.csr_reference(ei, n)
    assert torch.equal(g.rowptr.cpu().long(), ref["rowptr"])
    assert torch.equal(g.csr_eid.cpu().long(), ref["eid"])
    assert torch.equal(g.csr_src.cpu().long(), ref["src"])
    assert torch.equal(g.colptr.cpu().long(), ref["colptr"])
    pos = torch.empty(ei.shape[1], dtype=torch.long)
    pos[ref["eid"]] = torch.arange(ei.shape[1])
    assert torch.equal(g.pos_of_eid.cpu().long(), pos)
    assert torch.equal(g.csc_pos.cpu().long(), pos[ref["eid_by_src"]])
    assert torch.equal(g.csc_dst.cpu().long(), ref["dst_by_src"])
    assert torch.equal(g.in_degree.cpu().long(), ref["in_degree"])
    return g


def test_csr_molecular_batch_bit_exact():
    _, b = make_batch(256, seed=11)
    _check_csr(b["edge_index"], b["x"].shape[0])


@pytest.mark.parametrize("n,e,seed", [(1, 0, 0), (7, 0, 1), (5, 40, 2), (1000, 5000, 3), (300, 40000, 4), (70000, 150000, 5)])
def test_csr_random_multigraph_bit_exact(n, e, seed):
    """Random COO with duplicates, self loops, isolated nodes and hub rows (> 32 entries: heap-sort path)."""
    gen = torch.Generator().manual_seed(seed)
    ei = torch.randint(0, n, (2, e), generator=gen)
    if e > 100:
        ei[1, : e // 3] = 0                      # one hub destination
    _check_csr(ei, n)


def test_csr_rejects_out_of_range():
    ei = torch.tensor([[0, 1, 5], [1, 0, 2]])
    with pytest.raises(IndexError):
        graph.GraphIndex.build(ei.to(DEV), 3, check=True)


def test_degree_histogram_equals_calc_deg():
    graphs, b = make_batch(200, seed=5)
    g = graph.GraphIndex.build(b["edge_index"].to(DEV), b["x"].shape[0])
    hist, mx = graph.calc_deg_from_index(g)
    ref = O.calc_deg(oracle_graphs(graphs))
    assert hist[: int(mx.item()) + 1].tolist() == ref
    assert int(hist.sum()) == b["x"].shape[0]


def test_graph_ptr_from_batch():
    _, b = make_batch(100, seed=6)
    gp = ops.graph_ptr_from_batch(b["batch"].to(DEV), 100).cpu().long()
    ref = torch.zeros(101, dtype=torch.long)
    ref[1:] = torch.cumsum(torch.bincount(b["batch"], minlength=100), 0)
    assert torch.equal(gp, ref)


def test_packed_store_collate_equals_pyg_collate():
    graphs = synth.make_graphs(300, 21, 1, 30, 0.1)
    store = graph.DeviceGraphStore(synth.pack(graphs), DEV)
    ids = np.random.default_rng(0).permutation(300)[:128]
    b = store.collate(ids)
    ref = O.collate([oracle_graphs(graphs)[i] for i in ids])
    assert torch.equal(b.x.cpu().long(), ref.x)
    assert torch.equal(torch.stack([b._src32, b._dst32]).cpu().long(), ref.edge_index)
    assert torch.equal(b.edge_attr.cpu().long(), ref.edge_attr)
    assert torch.equal(b.batch.cpu().long(), ref.batch)
    assert torch.equal(b.graph_ptr.cpu().long(), ref.ptr)
    assert torch.equal(b.para.cpu(), ref.para)
    csr = O.csr_reference(ref.edge_index, ref.num_nodes)
    assert torch.equal(b.index.csr_eid.cpu().long(), csr["eid"])
    assert store.calc_deg(100) == O.calc_deg(oracle_graphs(graphs))


def test_edge_codes_and_narrow():
    _, b = make_batch(64, seed=8)
    ea = b["edge_attr"]
    g = graph.GraphIndex.build(b["edge_index"].to(DEV), b["x"].shape[0])
    ea8 = ops.narrow_u8(ea.to(DEV))
    assert torch.equal(ea8.cpu().long(), ea)
    codes = ops.edge_codes(ea8, g.csr_eid).cpu().long()
    sorted_ea = ea[g.csr_eid.cpu().long()]
    assert torch.equal(codes, (sorted_ea[:, 0] * 6 + sorted_ea[:, 1]) * 2 + sorted_ea[:, 2])


def test_host_batch_prefetcher_delivers_every_batch_in_order():
    """graph.HostBatchPrefetcher: pinned host tuples -> device tuples, copy of batch i+1 issued ahead on a copy stream;
    contents and order must be exact, non-tensor members pass through, unpinned input is rejected."""
    gen = torch.Generator().manual_seed(3)
    host = [(torch.randint(0, 1 << 40, (1000 + 37 * i, 9), generator=gen).pin_memory(),
             torch.randn(513 * (i + 1), generator=gen).pin_memory(), i) for i in range(7)]
    got = list(graph.HostBatchPrefetcher(iter(host), torch.device("cuda", 0)))
    assert len(got) == len(host)
    for (a, b, k), (ha, hb, hk) in zip(got, host):
        s = (a.double().sum() + b.double().sum()).item()          # consume on the current stream
        assert a.is_cuda and b.is_cuda and k == hk and torch.equal(a.cpu(), ha) and torch.equal(b.cpu(), hb)
        want = (ha.double().sum() + hb.double().sum()).item()
        assert abs(s - want) <= 1e-9 * abs(want)
    with pytest.raises(ValueError):
        graph.HostBatchPrefetcher(iter([(torch.zeros(4),)]), torch.device("cuda", 0))


def test_host_batch_prefetcher_prepare_next_builds_the_index_ahead():
    """prepare_next(fn) runs fn on the batch the NEXT __next__ returns (after its copy has landed on the current stream):
    the memoised CSR of that batch must be the one forward() then finds, and be bit-exact."""
    _, b1 = make_batch(24, seed=4)
    _, b2 = make_batch(16, seed=5)
    host = [(b["x"].pin_memory(), b["edge_index"].pin_memory()) for b in (b1, b2)]
    feed = graph.HostBatchPrefetcher(iter(host), torch.device("cuda", 0))
    seen = []

    def index_ahead(t):
        seen.append((t[1].data_ptr(), graph.graph_index_for(t[1], int(t[0].shape[0]))))

    feed.prepare_next(index_ahead)                     # batch 0, before it is handed out
    x0, ei0 = next(feed)
    assert seen[0][0] == ei0.data_ptr() and graph.graph_index_for(ei0, x0.shape[0]) is seen[0][1]
    feed.prepare_next(index_ahead)                     # batch 1 while batch 0 is "in use"
    x1, ei1 = next(feed)
    g1 = graph.graph_index_for(ei1, x1.shape[0])
    assert g1 is seen[1][1]
    ref = O.csr_reference(b2["edge_index"], b2["x"].shape[0])
    assert torch.equal(g1.csr_eid.cpu().long(), ref["eid"]) and torch.equal(g1.rowptr.cpu().long(), ref["rowptr"])
    feed.prepare_next(index_ahead)                     # nothing staged any more: a no-op
    assert len(seen) == 2
    with pytest.raises(StopIteration):
        next(feed)


# ------------------------------------------------------------------ invalid inputs (graph.py:66)
def _small_model(cfg_base=None):
    import copy
    import gnnepcsaft_b200 as G
    from tests.util import PNA_CFG1, deg_hist
    graphs, b = make_batch(8, seed=11)
    cfg = dict(cfg_base or PNA_CFG1, deg=deg_hist(graphs))
    torch.manual_seed(0)
    return G.GNNePCSAFT(copy.deepcopy(cfg)).to(DEV).eval(), b


@pytest.mark.parametrize("what", ["edge", "atom", "bond"])
def test_out_of_range_inputs_raise_like_pyg_and_never_touch_foreign_memory(what):
    """PyG raises on an edge_index entry >= num_nodes and torch.nn.Embedding on an id outside its vocabulary.  Here the
    flags ride the host reads the path pays anyway: ``int(batch[-1]) + 1`` of the reference-signature forward(), the
    max-degree readback of the PNA path when the model has no degree hint, and otherwise the lagged status copy that the
    next forward checks.  A path without
    any host read (GATv2 with num_graphs=) skips / zeroes the offending entries -- finite output, nothing uninitialised
    is ever used as an index -- and Batch.validate() / GraphIndex.check() report them."""
    from gnnepcsaft_b200 import graph as gr
    from tests.util import GAT_CFG_SMALL
    model, b = _small_model()
    gat, _ = _small_model(GAT_CFG_SMALL)
    x, ei, ea, bt = (b[k].clone() for k in ("x", "edge_index", "edge_attr", "batch"))
    G = int(bt[-1]) + 1
    if what == "edge":
        ei[1, 7] = x.shape[0] + 5
        ei[0, 3] = -2
    elif what == "atom":
        x[4, 0] = 119                                     # atomic-number vocabulary is 0..118
        x[9, 8] = 7
    else:
        ea[5, 1] = 6                                      # bond stereo vocabulary is 0..5
    args = [t.to(DEV) for t in (x, ei, ea, bt)]
    with torch.no_grad():
        for m in (model, gat):
            gr._INDEX_CACHE.clear()
            with pytest.raises(IndexError):
                m(*args)                                  # reference signature: the one host read checks the flags
        # PNA with num_graphs=: the dataset histogram bounds the in-degree (model.degree_hint), so nothing is read back
        # during the call -- memory-safe, finite -- and the batch's status words, copied to pinned memory on a side
        # stream, make the NEXT forward raise; without the hint the max-degree read-back raises at once
        gr._INDEX_CACHE.clear()
        out = model(*args, num_graphs=G)
        assert bool(torch.isfinite(out).all())
        gr._INDEX_CACHE.clear()
        with pytest.raises(IndexError):
            model(*args, num_graphs=G)
        hint, model.degree_hint = model.degree_hint, None
        model._hinted = None
        gr._INDEX_CACHE.clear()
        with pytest.raises(IndexError):
            model(*args, num_graphs=G)
        model.degree_hint = hint
        gr._INDEX_CACHE.clear()
        out = gat(*args, num_graphs=G)                    # sync-free path: memory-safe, flags left on the device
        assert out.shape == (G, 2) and bool(torch.isfinite(out).all())
        gr._INDEX_CACHE.clear()
        bb = gr.batch_from_reference_tensors(*args, num_graphs=G)
        with pytest.raises(IndexError):
            bb.validate()
        if what == "edge":
            g = bb.index
            E = ei.shape[1]
            nvalid = int(g.rowptr[-1])
            assert nvalid == E - 2                        # two edges skipped; the slot tails are zero-filled
            for t in (g.csr_src, g.csr_eid, g.csc_pos, g.csc_dst):
                assert bool((t[nvalid:] == 0).all()) and int(t.min()) >= 0 and int(t.max()) < max(E, x.shape[0])
    # valid input: no flag, no exception
    good = [b[k].to(DEV) for k in ("x", "edge_index", "edge_attr", "batch")]
    gr._INDEX_CACHE.clear()
    with torch.no_grad():
        model(*good)
        gat(*good, num_graphs=G)
        gr.batch_from_reference_tensors(*good, num_graphs=G).validate()


def test_packed_store_is_validated_on_the_host():
    from gnnepcsaft_b200 import graph as gr, synth
    graphs, _ = make_batch(6, seed=12)
    p = synth.pack(graphs)
    gr.DeviceGraphStore(p, DEV)
    bad = synth.pack(graphs)
    bad.x[3, 0] = 200
    with pytest.raises(IndexError):
        gr.DeviceGraphStore(bad, DEV)
    bad = synth.pack(graphs)
    bad.dst[2] = 120
    with pytest.raises(IndexError):
        gr.DeviceGraphStore(bad, DEV)


def test_reference_file_to_device_batches(tmp_path):
    """SURVEY 8(f) rank 2 end to end: a processed dataset file in the reference's (data, slices) layout -> packed device
    store -> shuffled epoch of device-collated batches, each bit-exactly the PyG collation of the graphs the DataLoader
    would have put into it."""
    from gnnepcsaft_b200 import dataset
    graphs = synth.make_graphs(70, 33, 1, 30, 0.1)
    p = tmp_path / "es_graph_data.pt"
    torch.save(dataset.collate_graphs(graphs), p)
    store = graph.DeviceGraphStore(dataset.pack_collated(*dataset.load_collated(p)), DEV)
    assert store.num_graphs == 70 and store.calc_deg() == O.calc_deg(oracle_graphs(graphs))
    torch.manual_seed(9)
    want_order = [b.tolist() for b in torch.utils.data.DataLoader(torch.arange(70), batch_size=16, shuffle=True)]
    torch.manual_seed(9)
    loader = dataset.EpochLoader(store, 16, shuffle=True)
    og = oracle_graphs(graphs)
    seen = []
    for b, ids in zip(loader, want_order):
        ref = O.collate([og[i] for i in ids])
        assert b.num_graphs == len(ids)
        assert torch.equal(b.x.cpu().long(), ref.x) and torch.equal(b.edge_attr.cpu().long(), ref.edge_attr)
        assert torch.equal(torch.stack([b._src32, b._dst32]).cpu().long(), ref.edge_index)
        assert torch.equal(b.batch.cpu().long(), ref.batch) and torch.equal(b.para.cpu(), ref.para)
        seen += ids
    assert sorted(seen) == list(range(70)) and len(loader) == 5


def test_packed_store_is_validated_on_the_device_and_reports_what_numpy_would():
    """DeviceGraphStore uploads, then checks on the device in molecule chunks: a bad endpoint / atom id / bond id raises
    the IndexError the host-side check used to raise, and the largest in-degree equals the numpy bincount -- also when
    the chunk size forces many chunks."""
    import copy
    from gnnepcsaft_b200 import synth
    packed = synth.pack(synth.make_graphs(300, 11, 5, 40, 0.0))
    store = graph.DeviceGraphStore(packed, torch.device(DEV))
    deg = np.bincount(np.repeat(packed.node_ptr[:-1], np.diff(packed.edge_ptr)) + packed.dst.astype(np.int64),
                      minlength=int(packed.node_ptr[-1]))
    assert store.max_degree == int(deg.max())
    assert store._validate_on_device(edges_per_chunk=97) == (False, False, False, int(deg.max()))
    for field, value, what in (("dst", 999, "edge endpoint"), ("src", -1, "edge endpoint"), ("x", 200, "atom feature"),
                               ("edge_attr", 77, "bond feature")):
        bad = copy.deepcopy(packed)
        arr = getattr(bad, field)
        arr.reshape(-1)[arr.size // 2] = value
        with pytest.raises(IndexError, match=what):
            graph.DeviceGraphStore(bad, torch.device(DEV))
