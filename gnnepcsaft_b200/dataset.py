"""Dataset front-end of the device-resident store: the reference's on-disk format in, shuffled
device-collated mini-batches out.

* ``load_collated(path)`` reads what the reference's datasets write and load -- ``torch.save(self.collate(datalist), path)``
  / ``self.data, self.slices = torch.load(path, weights_only=False)`` (gnnepcsaft/data/graphdataset.py:44, 89, 120, 152,
  181-183, 220): a ``(data, slices)`` tuple where ``data`` is a PyG ``Data`` holding every graph's tensors concatenated
  (``x`` along dim 0, ``edge_index`` along dim 1 with GRAPH-LOCAL node ids -- ``InMemoryDataset.collate`` does not offset
  them --, ``edge_attr`` / ``para`` / ``assoc`` / ``munanb`` along dim 0) and ``slices[key]`` the int64 ``[G+1]`` boundaries
  along that dim.  ``torch_geometric`` is NOT needed to read the file: its classes are unpickled as attribute bags.
* ``pack_collated(data, slices)`` turns that into the compact ``synth.PackedGraphs`` (uint8 features, int16 endpoints)
  that ``graph.DeviceGraphStore`` uploads once.
* ``EpochLoader`` replaces ``DataLoader(train_dataset, batch_size=B, shuffle=True, num_workers=os.cpu_count())``
  (gnnepcsaft/train/train.py:59-64): one pass over the store per ``iter()``, in the SAME order torch's ``RandomSampler``
  draws from the global RNG (so a run seeded like the reference's sees the same mini-batches), every batch collated on
  the device (``DeviceGraphStore.collate``: block-diagonal batch + CSR, no CPU workers, no int64 H2D).  Under data
  parallelism each rank takes a contiguous slice of every global batch (``train.shard_range``).
"""

from __future__ import annotations

import io
import pickle
from typing import Dict, Iterator, Mapping, Optional, Tuple

import numpy as np
import torch
from torch import Tensor

from .synth import PackedGraphs

NODE_KEY, EDGE_KEY, EATTR_KEY = "x", "edge_index", "edge_attr"


# ------------------------------------------------------------------ reading the reference's .pt without torch_geometric
class _Bag:
    """Stand-in for any ``torch_geometric.*`` class met while unpickling: keeps whatever state the pickle carries."""

    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        if isinstance(state, dict):
            self.__dict__.update(state)
        elif isinstance(state, tuple) and len(state) == 2 and isinstance(state[1], dict):
            if isinstance(state[0], dict):
                self.__dict__.update(state[0])
            self.__dict__.update(state[1])
        else:
            self.__dict__["_state"] = state


class _Unpickler(pickle.Unpickler):
    def find_class(self, module, name):
        if module.split(".")[0] == "torch_geometric":
            try:
                return super().find_class(module, name)          # the real class when PyG is installed
            except Exception:  # noqa: BLE001
                return type(name, (_Bag,), {"__module__": module})
        return super().find_class(module, name)


class _PickleModule:
    """``pickle_module=`` for ``torch.load``: the stock pickle with the tolerant ``find_class`` above."""
    __name__ = "gnnx_pickle"
    Unpickler = _Unpickler
    load = staticmethod(lambda f, **kw: _Unpickler(f, **kw).load())
    loads = staticmethod(lambda b, **kw: _Unpickler(io.BytesIO(b), **kw).load())
    dump, dumps, Pickler = pickle.dump, pickle.dumps, pickle.Pickler
    HIGHEST_PROTOCOL, DEFAULT_PROTOCOL = pickle.HIGHEST_PROTOCOL, pickle.DEFAULT_PROTOCOL
    PickleError, PicklingError, UnpicklingError = pickle.PickleError, pickle.PicklingError, pickle.UnpicklingError


def _tensor_items(obj) -> Dict[str, object]:
    """Every (key, value) of a PyG ``Data`` (real or unpickled as a bag) or of a plain mapping."""
    if isinstance(obj, Mapping):
        return dict(obj)
    if hasattr(obj, "to_dict"):                                   # a real torch_geometric Data
        return dict(obj.to_dict())
    d = getattr(obj, "__dict__", {})
    store = d.get("_store", None)
    if store is not None:                                         # Data -> GlobalStorage -> _mapping
        sd = getattr(store, "__dict__", {})
        if "_mapping" in sd:
            return dict(sd["_mapping"])
        return {k: v for k, v in sd.items() if not k.startswith("_")}
    return {k: v for k, v in d.items() if not k.startswith("_")}


def load_collated(path) -> Tuple[Dict[str, object], Dict[str, Tensor]]:
    """``(data, slices)`` of a processed reference dataset file as two plain dicts.  Accepts the reference's
    ``torch.save(self.collate(datalist))`` tuple, PyG >= 2.4's ``(data.to_dict(), slices, data_cls)`` triple, and a
    tuple whose first element already is a dict (the synthetic fixtures of ``tests/``)."""
    obj = torch.load(path, map_location="cpu", weights_only=False, pickle_module=_PickleModule)
    if not isinstance(obj, (tuple, list)) or len(obj) < 2:
        raise ValueError(f"{path}: expected the (data, slices) tuple InMemoryDataset.collate() returns, got {type(obj).__name__}")
    data, slices = _tensor_items(obj[0]), _tensor_items(obj[1])
    for k in (NODE_KEY, EDGE_KEY, EATTR_KEY):
        if k not in data or k not in slices:
            raise ValueError(f"{path}: the collated dataset has no '{k}' (keys: {sorted(data)})")
    return data, {k: torch.as_tensor(v, dtype=torch.int64) for k, v in slices.items()}


# ------------------------------------------------------------------ (data, slices) -> packed store
def pack_collated(data: Mapping[str, object], slices: Mapping[str, Tensor]) -> PackedGraphs:
    """Compact the collated int64 layout (72 B per atom, 40 B per directed edge) into the packed store (9 B / 7 B)."""
    x = torch.as_tensor(data[NODE_KEY])
    ei = torch.as_tensor(data[EDGE_KEY])
    ea = torch.as_tensor(data[EATTR_KEY])
    node_ptr = torch.as_tensor(slices[NODE_KEY], dtype=torch.int64).numpy().copy()
    edge_ptr = torch.as_tensor(slices[EDGE_KEY], dtype=torch.int64).numpy().copy()
    G = node_ptr.shape[0] - 1
    if x.dim() != 2 or x.shape[1] != 9 or ei.dim() != 2 or ei.shape[0] != 2 or ea.shape != (ei.shape[1], 3):
        raise ValueError(f"pack_collated: unexpected shapes x {tuple(x.shape)}, edge_index {tuple(ei.shape)}, edge_attr {tuple(ea.shape)}")
    if edge_ptr.shape[0] != G + 1 or node_ptr[0] != 0 or edge_ptr[0] != 0 or node_ptr[-1] != x.shape[0] or edge_ptr[-1] != ei.shape[1]:
        raise ValueError("pack_collated: slices do not delimit the concatenated tensors")
    if (np.diff(node_ptr) < 0).any() or (np.diff(edge_ptr) < 0).any():
        raise ValueError("pack_collated: slices must be non-decreasing")
    if not np.array_equal(torch.as_tensor(slices[EATTR_KEY]).numpy(), edge_ptr):
        raise ValueError("pack_collated: edge_attr and edge_index are sliced differently")
    if G and int(np.diff(node_ptr).max()) > np.iinfo(np.int16).max:
        raise ValueError("pack_collated: a graph with more than 32767 nodes does not fit the int16 endpoints of the store")
    for name, t in ((NODE_KEY, x), (EATTR_KEY, ea), (EDGE_KEY, ei)):
        if t.numel() and (int(t.min()) < 0 or int(t.max()) > (255 if name != EDGE_KEY else 32767)):
            raise IndexError(f"pack_collated: '{name}' holds values outside the packed store's range")

    def target(key: str, width: int) -> np.ndarray:
        if key not in data:
            return np.zeros((G, width), np.float32)
        t = torch.as_tensor(data[key]).to(torch.float32).reshape(G, -1)
        if t.shape[1] != width:
            raise ValueError(f"pack_collated: '{key}' is [G, {t.shape[1]}], expected [G, {width}]")
        return t.numpy().copy()

    return PackedGraphs(
        x=x.to(torch.uint8).numpy().copy(), src=ei[0].to(torch.int16).numpy().copy(), dst=ei[1].to(torch.int16).numpy().copy(),
        edge_attr=ea.to(torch.uint8).numpy().copy(), node_ptr=node_ptr, edge_ptr=edge_ptr,
        para=target("para", 3), assoc=target("assoc", 2))


def collate_graphs(graphs) -> Tuple[Dict[str, Tensor], Dict[str, Tensor]]:
    """What ``InMemoryDataset.collate(datalist)`` produces for a list of ``synth.MolGraph``: concatenated int64 tensors
    with graph-local ``edge_index`` and the ``slices`` dict.  Used to write synthetic files in the reference's layout."""
    t = torch.from_numpy
    n = np.array([g.num_nodes for g in graphs], np.int64)
    e = np.array([g.edge_index.shape[1] for g in graphs], np.int64)
    G = len(graphs)
    node_ptr = t(np.concatenate([[0], np.cumsum(n)]).astype(np.int64))
    edge_ptr = t(np.concatenate([[0], np.cumsum(e)]).astype(np.int64))
    per_graph = torch.arange(G + 1, dtype=torch.int64)
    data = {
        "x": t(np.concatenate([g.x for g in graphs]).astype(np.int64)),
        "edge_index": t(np.concatenate([g.edge_index for g in graphs], axis=1).astype(np.int64)),
        "edge_attr": t(np.concatenate([g.edge_attr for g in graphs]).astype(np.int64)),
        "para": t(np.concatenate([g.para for g in graphs]).astype(np.float32)),
        "assoc": t(np.concatenate([g.assoc for g in graphs]).astype(np.float32)),
    }
    slices = {"x": node_ptr, "edge_index": edge_ptr, "edge_attr": edge_ptr.clone(), "para": per_graph, "assoc": per_graph.clone()}
    return data, slices


# ------------------------------------------------------------------ shuffled epochs over the device-resident store
def random_sampler_order(n: int, generator: Optional[torch.Generator] = None, draw_base_seed: bool = True) -> np.ndarray:
    """The index order ``DataLoader(dataset_of_n, shuffle=True)`` visits in one epoch, reproduced draw for draw:
    the loader iterator first takes a base seed from the global (or given) generator, then ``RandomSampler`` seeds a
    fresh generator from the global RNG and yields ``torch.randperm(n, generator=that)``."""
    if draw_base_seed:
        torch.empty((), dtype=torch.int64).random_(generator=generator)
    if generator is None:
        seed = int(torch.empty((), dtype=torch.int64).random_().item())
        generator = torch.Generator()
        generator.manual_seed(seed)
    return torch.randperm(n, generator=generator).numpy()


class EpochLoader:
    """Mini-batches of a ``DeviceGraphStore`` -- ``len()`` / iteration semantics of ``DataLoader(batch_size=B, shuffle=S,
    drop_last=D)``; each ``iter()`` is one epoch.  ``rank`` / ``world``: this rank's slice of every global batch."""

    def __init__(self, store, batch_size: int, shuffle: bool = True, drop_last: bool = False,
                 generator: Optional[torch.Generator] = None, rank: int = 0, world: int = 1):
        if batch_size <= 0:
            raise ValueError("EpochLoader: batch_size must be positive")
        self.store, self.batch_size, self.shuffle, self.drop_last = store, int(batch_size), shuffle, drop_last
        self.generator, self.rank, self.world = generator, rank, world
        self.last_order: Optional[np.ndarray] = None

    def __len__(self) -> int:
        n, b = self.store.num_graphs, self.batch_size
        return n // b if self.drop_last else (n + b - 1) // b

    def __iter__(self) -> Iterator:
        from .train import shard_range
        n, b = self.store.num_graphs, self.batch_size
        order = random_sampler_order(n, self.generator) if self.shuffle else np.arange(n, dtype=np.int64)
        self.last_order = order
        for k in range(len(self)):
            ids = order[k * b:(k + 1) * b]
            if self.world > 1:
                lo, hi = shard_range(len(ids), self.rank, self.world)
                ids = ids[lo:hi]
            yield self.store.collate(ids)
