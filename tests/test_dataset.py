"""Dataset front-end: the reference's processed ``(data, slices)`` file is read without
torch_geometric, packs to the same store as ``synth.pack``, and ``EpochLoader`` visits the graphs in the order torch's
``DataLoader(shuffle=True)`` would (reference: data/graphdataset.py:181-183, 220; train/train.py:59-64)."""
import sys
import types

import numpy as np
import pytest
import torch

from gnnepcsaft_b200 import dataset, synth


def _graphs(n=40, seed=5):
    return synth.make_graphs(n, seed, 1, 30, 0.1)


def _assert_same_pack(a, b):
    for f in ("x", "src", "dst", "edge_attr", "node_ptr", "edge_ptr", "para", "assoc"):
        assert np.array_equal(getattr(a, f), getattr(b, f)), f


def test_dict_file_roundtrip(tmp_path):
    graphs = _graphs()
    data, slices = dataset.collate_graphs(graphs)
    # InMemoryDataset.collate keeps edge_index graph-local (no node offset)
    assert int(data["edge_index"].max()) < max(g.num_nodes for g in graphs)
    p = tmp_path / "es_graph_data.pt"
    torch.save((data, slices), p)
    d2, s2 = dataset.load_collated(p)
    _assert_same_pack(dataset.pack_collated(d2, s2), synth.pack(graphs))


def test_pyg_pickle_is_read_without_torch_geometric(tmp_path):
    """The reference writes ``torch.save(self.collate(datalist), path)``: a pickled torch_geometric ``Data`` whose
    ``_store`` (a ``GlobalStorage``) keeps the tensors in ``_mapping``.  Build such a pickle with stand-in classes living
    in a fake ``torch_geometric`` module tree, drop the modules again, and read the file with nothing of PyG importable."""
    graphs = _graphs(17, 9)
    data, slices = dataset.collate_graphs(graphs)
    names = ["torch_geometric", "torch_geometric.data", "torch_geometric.data.data", "torch_geometric.data.storage"]
    assert not any(n in sys.modules for n in names), "this test needs PyG to be absent"
    mods = {n: types.ModuleType(n) for n in names}

    class GlobalStorage:
        def __init__(self, mapping):
            self._mapping = dict(mapping)
            self._parent = None

    class Data:
        def __init__(self, mapping):
            self._edge_attr_cls = None
            self._tensor_attr_cls = None
            self._store = GlobalStorage(mapping)

    GlobalStorage.__module__, GlobalStorage.__qualname__ = "torch_geometric.data.storage", "GlobalStorage"
    Data.__module__, Data.__qualname__ = "torch_geometric.data.data", "Data"
    mods["torch_geometric.data.storage"].GlobalStorage = GlobalStorage
    mods["torch_geometric.data.data"].Data = Data
    p = tmp_path / "tml_graph_data.pt"
    sys.modules.update(mods)
    try:
        torch.save((Data(dict(data, InChI=["x"] * len(graphs))), slices), p)
    finally:
        for n in names:
            sys.modules.pop(n, None)
    with pytest.raises(Exception):
        torch.load(p, weights_only=False)                        # the stock loader needs torch_geometric
    d2, s2 = dataset.load_collated(p)
    assert d2["InChI"] == ["x"] * len(graphs)
    _assert_same_pack(dataset.pack_collated(d2, s2), synth.pack(graphs))


def test_pack_collated_rejects_bad_input():
    data, slices = dataset.collate_graphs(_graphs(6, 2))
    bad = dict(slices, edge_attr=slices["edge_attr"].clone())
    bad["edge_attr"][1] += 1
    with pytest.raises(ValueError):
        dataset.pack_collated(data, bad)
    with pytest.raises(IndexError):
        dataset.pack_collated(dict(data, x=data["x"] + 300), slices)
    with pytest.raises(ValueError):
        dataset.pack_collated(dict(data, x=data["x"][:-1]), slices)


def test_missing_targets_become_zeros():
    data, slices = dataset.collate_graphs(_graphs(5, 3))
    data.pop("assoc")
    pk = dataset.pack_collated(data, slices)
    assert pk.assoc.shape == (5, 2) and not pk.assoc.any() and pk.para.any()


@pytest.mark.parametrize("n,bs,drop_last", [(103, 16, False), (103, 16, True), (64, 64, False), (5, 8, False)])
def test_epoch_order_is_the_dataloaders(n, bs, drop_last):
    """Same seed, same mini-batches as ``torch.utils.data.DataLoader(range(n), batch_size, shuffle=True)`` -- the PyG
    loader of train/train.py:59-64 only swaps the collate function."""

    class Store:                                         # host stand-in: collate returns the ids it was asked for
        num_graphs = n

        def collate(self, ids):
            return np.asarray(ids).copy()

    for epoch_seed in (0, 123):
        torch.manual_seed(epoch_seed)
        want = [b.numpy() for b in torch.utils.data.DataLoader(torch.arange(n), batch_size=bs, shuffle=True, drop_last=drop_last)]
        want2 = [b.numpy() for b in torch.utils.data.DataLoader(torch.arange(n), batch_size=bs, shuffle=True, drop_last=drop_last)]
        torch.manual_seed(epoch_seed)
        loader = dataset.EpochLoader(Store(), bs, shuffle=True, drop_last=drop_last)
        got, got2 = list(loader), list(loader)               # two epochs: the RNG stream advances like the reference's
        assert len(loader) == len(want) == len(got)
        for a, b in zip(got + got2, want + want2):
            assert np.array_equal(a, b)
    # sequential + sharded: every rank a contiguous slice of each global batch, together the whole batch
    parts = [list(dataset.EpochLoader(Store(), bs, shuffle=False, drop_last=drop_last, rank=r, world=3)) for r in range(3)]
    whole = list(dataset.EpochLoader(Store(), bs, shuffle=False, drop_last=drop_last))
    for k, w in enumerate(whole):
        assert np.array_equal(np.concatenate([parts[r][k] for r in range(3)]), w)
