"""GPU batch collator: dst-sorted CSR, degree histograms, device-resident packed datasets.

Replaces, for the hot path, PyG's ``DataLoader`` / ``Batch.from_data_list`` (reference:
gnnepcsaft/train/train.py:59-75) and ``calc_deg`` (gnnepcsaft/train/utils.py:37-60).  All integer
results are bit-exact with the PyG ordering: the CSR permutation equals
``argsort(edge_index[1], stable=True)`` and the histogram equals ``bincount(degree(...))``.
"""

from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch
from torch import Tensor

from . import ops

ATOM_FEATURE_DIMS = (119, 5, 12, 12, 10, 6, 6, 2, 2)   # gnnepcsaft/data/ogb_utils.py:9-23
BOND_FEATURE_DIMS = (5, 6, 2)                           # gnnepcsaft/data/ogb_utils.py:24-33


class DegreeBuckets:
    """Nodes of one batch bucketed by in-degree (``ops.degree_buckets``): the row order of the degree-grouped
    first post_nn Linear of PNAConv.  Built once per batch, shared by all layers and by the backward."""

    __slots__ = ("n_groups", "node_pos", "row_node", "bucket_ptr", "err_flag", "capacity", "_tables")

    def __init__(self, rowptr: Tensor, n_groups: int, err: Optional[Tensor] = None):
        self.n_groups = n_groups
        self.node_pos, self.row_node, self.bucket_ptr, self.err_flag, self.capacity = ops.degree_buckets(rowptr, n_groups, err)
        self._tables = {}

    def split_table(self, splits: int) -> Tensor:
        t = self._tables.get(splits)
        if t is None:
            t = self._tables[splits] = ops.group_split_table(self.bucket_ptr, self.n_groups, splits)
        return t


class GraphIndex:
    """dst-sorted CSR + src-sorted view of one (batched, block-diagonal) ``edge_index``."""

    __slots__ = ("num_nodes", "num_edges", "rowptr", "csr_src", "csr_eid", "pos_of_eid", "colptr", "csc_pos",
                 "csc_dst", "err_flag", "_codes", "max_degree", "_buckets", "degree_hint", "_status_host")

    def __init__(self, num_nodes: int, num_edges: int, parts):
        self.num_nodes, self.num_edges = num_nodes, num_edges
        (self.rowptr, self.csr_src, self.csr_eid, self.pos_of_eid, self.colptr, self.csc_pos, self.csc_dst,
         self.err_flag) = parts
        self._codes = None
        self.max_degree = None        # host-side bound on the in-degree, when the collator knows one
        self.degree_hint = None       # a bound the CALLER vouches for (the dataset histogram behind PNAConv's `deg`): checked on the device
        self._buckets = None
        self._status_host = None      # (pinned copy of the status words, event) queued by the hinted bucketing

    def degree_buckets(self) -> Optional[DegreeBuckets]:
        """In-degree buckets of this batch, or ``None`` when it has more than ``ops.MAX_GROUPS`` degree classes
        (the caller then takes the general scaler path).  Without a host-side bound from the collator the maximum
        in-degree is read back once per batch (one device->host sync) -- together with the batch's error flags, so
        an out-of-range ``edge_index`` / feature id raises here like PyG / torch.nn.Embedding do, at no extra sync."""
        if self._buckets is None:
            if self.max_degree is None and self.degree_hint is not None and self.degree_hint < ops.MAX_GROUPS:
                # no read-back: bucket with the hinted number of degree classes; a node beyond it raises the batch's
                # ST_DEGREE_OVERFLOW status word (read_status() raises; GNNePCSAFT.forward reads it one call later)
                st = self.err_flag[ops.ST_DEGREE_OVERFLOW:ops.ST_DEGREE_OVERFLOW + 1]
                self._buckets = DegreeBuckets(self.rowptr, self.degree_hint + 1, st)
                # the status words travel to pinned memory on a copy stream, behind an event: reading them later
                # (read_status_async) waits for THIS point of the stream only, not for whatever was queued after it
                dev = self.rowptr.device
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(dev))
                cs = _status_stream(dev)
                cs.wait_event(ev)
                host = torch.empty(ops.STATUS_WORDS, dtype=torch.int32).pin_memory()
                with torch.cuda.stream(cs):
                    host.copy_(self.err_flag, non_blocking=True)
                    done = torch.cuda.Event()
                    done.record(cs)
                self.err_flag.record_stream(cs)
                self._status_host = (host, done)
                return self._buckets
            if self.max_degree is None:
                ops.degree_hist(self.rowptr, torch.zeros(1, dtype=torch.int64, device=self.rowptr.device),
                                self.err_flag[ops.ST_MAX_DEGREE:ops.ST_MAX_DEGREE + 1])
                self.max_degree = self.read_status()[ops.ST_MAX_DEGREE]
            self._buckets = DegreeBuckets(self.rowptr, self.max_degree + 1) if self.max_degree < ops.MAX_GROUPS else False
        return self._buckets or None

    def read_status_async(self) -> Optional[List[int]]:
        """The status words as they were right after a hinted ``degree_buckets()`` (copied to pinned memory on a side
        stream at that point): waits for that copy only.  ``None`` when no such copy was queued.  Raises like ``read_status``."""
        if self._status_host is None:
            return None
        host, done = self._status_host
        done.synchronize()
        return self._raise_on_flags(host.tolist())

    def read_status(self) -> List[int]:
        """The batch's status words (``ops.ST_*``) on the host -- ONE device->host sync; raises on any error flag."""
        return self._raise_on_flags(self.err_flag.tolist())

    @staticmethod
    def _raise_on_flags(vals: List[int]) -> List[int]:
        if vals[ops.ST_EDGE]:
            raise IndexError("edge_index contains node ids outside [0, num_nodes)")
        if vals[ops.ST_ATOM]:
            raise IndexError("x contains an atom feature id outside its OGB vocabulary (gnnepcsaft/data/ogb_utils.py:9-23)")
        if vals[ops.ST_BOND]:
            raise IndexError("edge_attr contains a bond feature id outside its OGB vocabulary (gnnepcsaft/data/ogb_utils.py:24-33)")
        if vals[ops.ST_DEGREE_OVERFLOW]:
            raise IndexError("a node's in-degree exceeds the degree histogram the PNA model was built with (`deg`): the degree-grouped "
                             "path was bucketed with too few classes and this batch's result is wrong; set model.degree_hint = None "
                             "(one read-back per batch) or rebuild the histogram with calc_deg over this data")
        return vals

    @staticmethod
    def build(edge_index: Tensor, num_nodes: int, check: bool = False) -> "GraphIndex":
        g = GraphIndex(num_nodes, edge_index.shape[1], ops.csr_build(edge_index, num_nodes))
        if check:
            g.check()
        return g

    @staticmethod
    def build_i32(src: Tensor, dst: Tensor, num_nodes: int, check: bool = False) -> "GraphIndex":
        g = GraphIndex(num_nodes, src.numel(), ops.csr_build(None, num_nodes, src, dst))
        if check:
            g.check()
        return g

    def check(self) -> None:
        """Raise like PyG / torch do on an out-of-range endpoint or feature id (costs one device->host sync)."""
        self.read_status()

    @property
    def in_degree(self) -> Tensor:
        return self.rowptr[1:] - self.rowptr[:-1]


_INDEX_CACHE: dict = {}
_STATUS_STREAMS: dict = {}


def _status_stream(device) -> "torch.cuda.Stream":
    key = device.index if device.index is not None else torch.cuda.current_device()
    s = _STATUS_STREAMS.get(key)
    if s is None:
        s = _STATUS_STREAMS[key] = torch.cuda.Stream(torch.device("cuda", key))
    return s


def graph_index_for(edge_index: Tensor, num_nodes: int) -> GraphIndex:
    """CSR for ``edge_index``, memoised on the tensor's identity so the L conv layers of one
    forward (which all receive the same ``edge_index`` tensor, models.py:208-216) sort once."""
    key = (edge_index.data_ptr(), edge_index._version, tuple(edge_index.shape), num_nodes, edge_index.device)
    hit = _INDEX_CACHE.get("k")
    if hit is not None and hit[0] == key:
        return hit[1]
    g = GraphIndex.build(edge_index, num_nodes)
    _INDEX_CACHE["k"] = (key, g, edge_index)      # keep the tensor alive so the pointer stays unique
    return g


def calc_deg_from_index(g: GraphIndex, nbins: int = 64):
    """In-degree histogram of one batch as device tensors ``(hist[int64 nbins], max_degree[int32 1])``."""
    hist = torch.zeros(nbins, dtype=torch.int64, device=g.rowptr.device)
    mx = torch.zeros(1, dtype=torch.int32, device=g.rowptr.device)
    ops.degree_hist(g.rowptr, hist, mx)
    return hist, mx


@dataclass
class Batch:
    """What ``GNNePCSAFT.forward`` / ``pred_with_bounds`` receive: the reference's ``Data`` fields
    plus the prebuilt device-side index structures (so nothing is re-derived per layer)."""
    x: Tensor                      # int64 [N,9] (reference layout) or uint8 [N,9]
    edge_index: Optional[Tensor]   # int64 [2,E] or None when built from the packed store
    edge_attr: Tensor              # int64 [E,3] or uint8 [E,3]
    batch: Optional[Tensor]        # int64 [N] or int32 [N]
    num_graphs: int = 0
    index: Optional[GraphIndex] = None
    graph_ptr: Optional[Tensor] = None
    para: Optional[Tensor] = None
    assoc: Optional[Tensor] = None

    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])

    def validate(self) -> None:
        """Raise ``IndexError`` if any endpoint / feature id of this batch was out of range (one device->host sync).
        The kernels are memory-safe either way (such ids are skipped / read as 0); paths that already synchronise
        (``forward(..., batch)`` without ``num_graphs``, the max-degree readback) check the same flags for free."""
        if self.index is not None:
            self.index.read_status()


class DeviceGraphStore:
    """A whole dataset kept packed in HBM (uint8 features, int16 local endpoints) -- SURVEY §8(f) rank 2."""

    def __init__(self, packed, device):
        # Upload first, then validate ONCE on the device (batches collated from the store need no check of their own): for
        # a 1.25 M-molecule screening shard the same checks in numpy (an [E] repeat, three [E] / [N, 9] comparisons, an
        # [N]-bin bincount for the largest in-degree) took 2.3 s of host time against 0.3 s for the upload itself.
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
        self.x, self.src, self.dst, self.edge_attr = t(packed.x), t(packed.src), t(packed.dst), t(packed.edge_attr)
        self.node_ptr, self.edge_ptr = t(packed.node_ptr), t(packed.edge_ptr)
        self.para, self.assoc = t(packed.para), t(packed.assoc)
        self.node_ptr_host, self.edge_ptr_host = packed.node_ptr, packed.edge_ptr   # numpy, for sizing on the host
        self.num_graphs = packed.num_graphs
        self.device = device
        bad_edge, bad_atom, bad_bond, max_deg = self._validate_on_device()
        if bad_edge:
            raise IndexError("packed store: an edge endpoint lies outside its molecule's node range")
        if bad_atom:
            raise IndexError("packed store: an atom feature id is outside its OGB vocabulary")
        if bad_bond:
            raise IndexError("packed store: a bond feature id is outside its OGB vocabulary")
        # largest in-degree of the dataset: every batch buckets its nodes into the same degree classes
        self.max_degree = max_deg

    def _validate_on_device(self, edges_per_chunk: int = 1 << 24):
        """Endpoint / vocabulary range checks and the largest in-degree, molecule chunks of <= ``edges_per_chunk`` edges at
        a time (bounded temporaries), ONE host read at the end -> (bad_edge, bad_atom, bad_bond, max_degree)."""
        dev = self.device
        flags = torch.zeros(4, dtype=torch.int64, device=dev)          # bad edge / atom / bond counts, max in-degree
        if self.x.numel():
            lim = torch.tensor(ATOM_FEATURE_DIMS, dtype=self.x.dtype, device=dev)
            flags[1] = (self.x >= lim).any().to(torch.int64)
        if self.edge_attr.numel():
            lim = torch.tensor(BOND_FEATURE_DIMS, dtype=self.edge_attr.dtype, device=dev)
            flags[2] = (self.edge_attr >= lim).any().to(torch.int64)
        ep, G = self.edge_ptr_host, self.num_graphs
        g0 = 0
        while g0 < G and self.src.numel():
            g1 = int(np.searchsorted(ep, ep[g0] + edges_per_chunk, side="right")) - 1      # last molecule that still fits
            g1 = min(G, max(g1, g0 + 1))
            e0, e1 = int(ep[g0]), int(ep[g1])
            n0, n1 = int(self.node_ptr_host[g0]), int(self.node_ptr_host[g1])
            if e1 > e0:
                ne = self.edge_ptr[g0 + 1:g1 + 1] - self.edge_ptr[g0:g1]
                nn = (self.node_ptr[g0 + 1:g1 + 1] - self.node_ptr[g0:g1]).to(torch.int32)
                n_of_edge = torch.repeat_interleave(nn, ne, output_size=e1 - e0)
                src, dst = self.src[e0:e1].to(torch.int32), self.dst[e0:e1].to(torch.int32)
                bad = (src < 0) | (dst < 0) | (src >= n_of_edge) | (dst >= n_of_edge)
                flags[0] += bad.any().to(torch.int64)
                off = torch.repeat_interleave((self.node_ptr[g0:g1] - n0).to(torch.int32), ne, output_size=e1 - e0)
                tgt = (off + dst).clamp_(0, max(n1 - n0 - 1, 0)).to(torch.int64)             # clamped: a bad endpoint is reported above
                deg = torch.bincount(tgt, minlength=max(n1 - n0, 1))
                flags[3] = torch.maximum(flags[3], deg.max())
            g0 = g1
        f = flags.tolist()
        return bool(f[0]), bool(f[1]), bool(f[2]), int(f[3])

    def collate(self, graph_ids: np.ndarray) -> Batch:
        """Block-diagonal batch of the selected molecules, built entirely on the device."""
        ids = np.asarray(graph_ids, dtype=np.int64)
        n = self.node_ptr_host[ids + 1] - self.node_ptr_host[ids]
        e = self.edge_ptr_host[ids + 1] - self.edge_ptr_host[ids]
        offs = np.zeros((3, ids.size + 1), np.int64)
        offs[0, :-1] = ids
        np.cumsum(n, out=offs[1, 1:])
        np.cumsum(e, out=offs[2, 1:])
        N, E = int(offs[1, -1]), int(offs[2, -1])
        dev_offs = torch.from_numpy(offs).to(self.device, non_blocking=True)
        gids = dev_offs[0, :-1].contiguous()
        x, src, dst, ea, batch = ops.collate_packed(self, gids, dev_offs[1], dev_offs[2], N, E)
        index = GraphIndex.build_i32(src, dst, N)
        index.max_degree = self.max_degree
        graph_ptr = dev_offs[1].to(torch.int32)
        b = Batch(x=x, edge_index=None, edge_attr=ea, batch=batch, num_graphs=int(ids.size), index=index,
                  graph_ptr=graph_ptr, para=self.para[gids], assoc=self.assoc[gids])
        b._src32, b._dst32 = src, dst
        return b

    def calc_deg(self, batch_graphs: int = 65536) -> List[int]:
        """Dataset in-degree histogram (reference ``calc_deg``), accumulated on the device batch by batch."""
        hist = torch.zeros(64, dtype=torch.int64, device=self.device)
        mx = torch.zeros(1, dtype=torch.int32, device=self.device)
        for lo in range(0, self.num_graphs, batch_graphs):
            b = self.collate(np.arange(lo, min(self.num_graphs, lo + batch_graphs)))
            ops.degree_hist(b.index.rowptr, hist, mx)
        m = int(mx.item())
        if m >= 63:
            raise ValueError("calc_deg: in-degree >= 63 not supported by the 64-bin histogram")
        return hist[: m + 1].tolist()


def batch_from_reference_tensors(x: Tensor, edge_index: Tensor, edge_attr: Tensor, batch: Optional[Tensor],
                                 num_graphs: Optional[int] = None) -> Batch:
    """Wrap the reference's int64 batch tensors (already on the device) and build the index once."""
    N = int(x.shape[0])
    index = graph_index_for(edge_index, N)
    st = index.err_flag
    # narrow the categorical features here, range-checked against the OGB vocabularies: the flags land in the batch's
    # status words, which the ONE host read below (or the max-degree readback, or Batch.validate()) checks
    if x.dtype != torch.uint8:
        x = ops.narrow_u8(x, ATOM_FEATURE_DIMS, st[ops.ST_ATOM:ops.ST_ATOM + 1])
    if edge_attr.dtype != torch.uint8:
        edge_attr = ops.narrow_u8(edge_attr, BOND_FEATURE_DIMS, st[ops.ST_BOND:ops.ST_BOND + 1])
    gp = None
    G = 1
    if batch is not None:
        if num_graphs is None:
            if N > 0:                                              # PyG syncs here too (int(batch.max())+1)
                st[ops.ST_LAST_GRAPH:ops.ST_LAST_GRAPH + 1].copy_(batch[-1:])
                num_graphs = index.read_status()[ops.ST_LAST_GRAPH] + 1
            else:
                num_graphs = 0
        G = num_graphs
        gp = ops.graph_ptr_from_batch(batch, G)
    return Batch(x=x, edge_index=edge_index, edge_attr=edge_attr, batch=batch, num_graphs=G, index=index, graph_ptr=gp)


class HostBatchPrefetcher:
    """Host -> device input pipeline for the reference-facing path (the reference's ``DataLoader(pin_memory=True)`` hands
    ``training_step`` one host batch at a time, train/train.py:59-75): iterates over tuples of PINNED host tensors and
    yields them as device tensors, with the copy of batch i+1 issued on a copy stream before batch i is returned, so
    the H2D transfer (16 MB of int64 at cfg2, ~0.6 ms) runs under the previous step's kernels.  Stream-ordered only:
    the consumer's stream waits on the copy's event, and the tensors are recorded on it for the caching allocator."""

    def __init__(self, host_batches, device):
        self.it = iter(host_batches)
        self.device = device
        self.copy_stream = torch.cuda.Stream(device)
        self._next = None
        self._stage()

    def _stage(self):
        try:
            host = next(self.it)
        except StopIteration:
            self._next = None
            return
        for t in host:
            if isinstance(t, Tensor) and not t.is_pinned():
                raise ValueError("HostBatchPrefetcher: host tensors must be pinned (the copy would not be asynchronous)")
        with torch.cuda.stream(self.copy_stream):
            dev = tuple(t.to(self.device, non_blocking=True) if isinstance(t, Tensor) else t for t in host)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self._next = (dev, ev)

    def __iter__(self):
        return self

    def _claim(self):
        """Make the staged batch usable on the current stream (idempotent)."""
        dev, ev = self._next
        if ev is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            for t in dev:
                if isinstance(t, Tensor):
                    t.record_stream(cur)
            self._next = (dev, None)
        return dev

    def prepare_next(self, fn) -> None:
        """Queue ``fn(device_tensors)`` for the batch the next ``__next__`` will return -- typically the parameter-free
        index work of that batch (``graph_index_for(edge_index, N).degree_buckets()``: CSR build, degree buckets and the
        one max-degree readback).  Called after a step's kernels are queued and BEFORE its result is read back, it keeps
        the device busy across the host's wait and leaves one host synchronisation per step instead of two."""
        if self._next is not None:
            fn(self._claim())

    def __next__(self):
        if self._next is None:
            raise StopIteration
        dev = self._claim()
        self._stage()
        return dev
