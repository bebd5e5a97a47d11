"""Deterministic synthetic OGB-featurised molecular graphs.

The reference's real inputs come from ``smiles2graph`` (gnnepcsaft/data/ogb_utils.py:93-147),
which needs rdkit; the benchmark configs in BASELINE.json are defined on *synthetic* graphs
with the same layout: ``x:int64[n,9]``, ``edge_index:int64[2,2b]`` with bond ``b`` emitted as
columns ``2b:(i,j)`` and ``2b+1:(j,i)`` (ogb_utils.py:125-129), ``edge_attr:int64[2b,3]``
identical for both directions, zero-bond graphs giving ``[2,0]`` / ``[0,3]`` (ogb_utils.py:137-139).
"""

from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

ATOM_FEATURE_DIMS = (119, 5, 12, 12, 10, 6, 6, 2, 2)   # data/ogb_utils.py:9-23
BOND_FEATURE_DIMS = (5, 6, 2)                           # data/ogb_utils.py:24-33
_ATOM_IDS = np.array([5, 6, 7, 8, 15, 16])              # C N O F S Cl (atomic number - 1)
_ATOM_P = np.array([0.70, 0.08, 0.12, 0.04, 0.03, 0.03])
_BOND_TYPE_P = np.array([0.62, 0.10, 0.03, 0.25])       # single / double / triple / aromatic
# bounds of the (m, sigma, epsilon/k) head and of the association head, train/models.py:167-172
PARA_LOWER = np.array([1.0, 1.9, 50.0])
PARA_UPPER = np.array([25.0, 4.5, 550.0])
ASSOC_LOWER = np.array([-math.log10(0.9), math.log10(200.0)])
ASSOC_UPPER = np.array([-math.log10(0.0001), math.log10(5000.0)])


@dataclass
class MolGraph:
    """One molecule in the reference's ``Data`` layout (all int64, like PyG)."""
    x: np.ndarray            # [n, 9]
    edge_index: np.ndarray   # [2, 2b]
    edge_attr: np.ndarray    # [2b, 3]
    para: np.ndarray         # [1, 3]   float32 target (m, sigma, epsilon/k)
    assoc: np.ndarray        # [1, 2]   float32 target (|log10 kappa_AB|, log10 epsilon_AB/k)

    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])


def _random_molecule_bonds(n: int, rng: np.random.Generator) -> List[Tuple[int, int]]:
    """Random tree with max degree 4 plus up to Binomial(3, 0.4) ring closures."""
    if n <= 1:
        return []
    deg = [0] * n
    adj = [set() for _ in range(n)]
    bonds: List[Tuple[int, int]] = []
    open_nodes = [0]                      # earlier nodes with degree < 4
    picks = rng.random(n)                 # one uniform per attached node
    for v in range(1, n):
        u = open_nodes[int(picks[v] * len(open_nodes))]
        bonds.append((u, v))
        adj[u].add(v); adj[v].add(u)
        deg[u] += 1; deg[v] += 1
        if deg[u] >= 4:
            open_nodes.remove(u)
        open_nodes.append(v)
    k = int(rng.binomial(3, 0.4))
    for _ in range(k):
        for _attempt in range(8):
            a, b = (int(t) for t in rng.integers(0, n, 2))
            if a != b and b not in adj[a] and deg[a] < 4 and deg[b] < 4:
                bonds.append((a, b))
                adj[a].add(b); adj[b].add(a)
                deg[a] += 1; deg[b] += 1
                break
    return bonds


def make_graph(n: int, rng: np.random.Generator) -> MolGraph:
    bonds = _random_molecule_bonds(n, rng)
    x = np.stack([rng.integers(0, d, n) for d in ATOM_FEATURE_DIMS], axis=1).astype(np.int64)
    x[:, 0] = _ATOM_IDS[rng.choice(len(_ATOM_IDS), n, p=_ATOM_P)]
    b = len(bonds)
    if b:
        bt = np.stack([rng.choice(4, b, p=_BOND_TYPE_P), np.zeros(b, np.int64), rng.integers(0, 2, b)], axis=1)
        ij = np.asarray(bonds, dtype=np.int64)
        edge_index = np.empty((2, 2 * b), np.int64)
        edge_index[0, 0::2], edge_index[1, 0::2] = ij[:, 0], ij[:, 1]
        edge_index[0, 1::2], edge_index[1, 1::2] = ij[:, 1], ij[:, 0]
        edge_attr = np.repeat(bt.astype(np.int64), 2, axis=0)
    else:
        edge_index = np.empty((2, 0), np.int64)
        edge_attr = np.empty((0, 3), np.int64)
    para = rng.uniform(PARA_LOWER, PARA_UPPER).astype(np.float32)[None]
    assoc = rng.uniform(ASSOC_LOWER, ASSOC_UPPER).astype(np.float32)[None]
    return MolGraph(x, edge_index, edge_attr, para, assoc)


def make_graphs(num_graphs: int, seed: int, n_min: int = 10, n_max: int = 40,
                single_atom_frac: float = 0.01) -> List[MolGraph]:
    """Training-style set: ``n ~ U{n_min..n_max}``, a fraction forced to one atom (zero edges)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    sizes = rng.integers(n_min, n_max + 1, num_graphs)
    if single_atom_frac > 0 and num_graphs > 1:
        forced = rng.random(num_graphs) < single_atom_frac
        sizes = np.where(forced, 1, sizes)
    return [make_graph(int(n), rng) for n in sizes]


def make_inference_graphs(num_graphs: int, seed: int) -> List[MolGraph]:
    """Screening-style set (BASELINE cfg5): ``n ~ U{1..100}``."""
    return make_graphs(num_graphs, seed, 1, 100, 0.0)


@dataclass
class PackedGraphs:
    """A set of molecules in one contiguous, compact store.

    Node / bond features as uint8, *graph-local* endpoints as int16, targets as float32;
    ``node_ptr`` / ``edge_ptr`` delimit graph ``g``.  ~9x smaller than the int64 PyG layout.
    """
    x: np.ndarray           # uint8  [Ntot, 9]
    src: np.ndarray         # int16  [Etot]   graph-local source
    dst: np.ndarray         # int16  [Etot]   graph-local destination
    edge_attr: np.ndarray   # uint8  [Etot, 3]
    node_ptr: np.ndarray    # int64  [G+1]
    edge_ptr: np.ndarray    # int64  [G+1]
    para: np.ndarray        # float32 [G, 3]
    assoc: np.ndarray       # float32 [G, 2]

    @property
    def num_graphs(self) -> int:
        return int(self.node_ptr.shape[0] - 1)


def pack(graphs: List[MolGraph]) -> PackedGraphs:
    n = np.array([g.num_nodes for g in graphs], np.int64)
    e = np.array([g.edge_index.shape[1] for g in graphs], np.int64)
    node_ptr = np.concatenate([[0], np.cumsum(n)])
    edge_ptr = np.concatenate([[0], np.cumsum(e)])
    ei = np.concatenate([g.edge_index for g in graphs], axis=1) if graphs else np.empty((2, 0), np.int64)
    return PackedGraphs(
        x=np.concatenate([g.x for g in graphs]).astype(np.uint8),
        src=ei[0].astype(np.int16), dst=ei[1].astype(np.int16),
        edge_attr=np.concatenate([g.edge_attr for g in graphs]).astype(np.uint8),
        node_ptr=node_ptr, edge_ptr=edge_ptr,
        para=np.concatenate([g.para for g in graphs]).astype(np.float32),
        assoc=np.concatenate([g.assoc for g in graphs]).astype(np.float32),
    )


def collate_numpy(graphs: List[MolGraph], lo: int = 0, hi: Optional[int] = None):
    """Host-side block-diagonal batch in the reference's int64 layout (what the PyG
    DataLoader hands to ``training_step``, train/train.py:59-75): returns
    ``x, edge_index, edge_attr, batch, para, assoc`` as numpy arrays."""
    gs = graphs[lo:hi]
    n = np.array([g.num_nodes for g in gs], np.int64)
    off = np.concatenate([[0], np.cumsum(n)])[:-1]
    x = np.concatenate([g.x for g in gs])
    edge_index = np.concatenate([g.edge_index + o for g, o in zip(gs, off)], axis=1)
    edge_attr = np.concatenate([g.edge_attr for g in gs])
    batch = np.repeat(np.arange(len(gs), dtype=np.int64), n)
    para = np.concatenate([g.para for g in gs])
    assoc = np.concatenate([g.assoc for g in gs])
    return x, edge_index, edge_attr, batch, para, assoc
