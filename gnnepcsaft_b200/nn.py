"""Drop-in modules for the PyG / ogb classes the reference model is built from.

Same constructor arguments, ``forward`` signatures, error behaviour and state-dict key names as
``torch_geometric.nn.{PNAConv, GATv2Conv, TransformerConv, BatchNorm}``, ``torch_geometric.nn.aggr.{Sum,Mean,Max}Aggregation``
and ``ogb.graphproppred.mol_encoder.{AtomEncoder, BondEncoder}`` as used at
gnnepcsaft/train/models.py:175-184, 441-495, 587-595 -- backed by the sm_100a kernels of libgnnx.
"""

from __future__ import annotations

import math
from typing import List, Optional, Sequence, Union

import torch
from torch import Tensor
from torch.nn import BatchNorm1d, Embedding, Linear, Module, ModuleList, Parameter, ReLU, Sequential

from . import functional as Fn
from . import ops
from .functional import EdgeTable
from .gat_fn import GATMeta, GATv2ConvFn
from .graph import ATOM_FEATURE_DIMS, BOND_FEATURE_DIMS, GraphIndex, graph_index_for
from .pna_fn import PNAConvFn, PNAMeta, grouped_ok, prepare as pna_prepare
from .tconv_fn import TConvMeta, TransformerConvFn

NUM_BOND_CODES = 60

_PNA_AGGREGATORS = ["mean", "min", "max", "std"]
_PNA_SCALERS = ["identity", "amplification", "attenuation"]


def _require_cuda(x: Tensor, who: str) -> None:
    if not x.is_cuda:
        raise RuntimeError(f"{who}: input is on {x.device}; the gnnx modules run on CUDA only (no CPU fallback)")


# ------------------------------------------------------------------ encoders (ogb)
class _FeatureEncoder(Module):
    list_name = ""
    dims: Sequence[int] = ()

    def __init__(self, emb_dim: int):
        super().__init__()
        embs = ModuleList()
        for d in self.dims:
            emb = Embedding(d, emb_dim)
            torch.nn.init.xavier_uniform_(emb.weight.data)
            embs.append(emb)
        setattr(self, self.list_name, embs)
        offs = [0]
        for d in self.dims:
            offs.append(offs[-1] + d)
        self.register_buffer("_offsets", torch.tensor(offs[:-1], dtype=torch.int32), persistent=False)

    def table(self) -> Tensor:
        return torch.cat([e.weight for e in getattr(self, self.list_name)], 0)

    def forward(self, x: Tensor, dropout_p: float = 0.0, seed: int = 0) -> Tensor:
        """``dropout_p`` > 0 (gnnx extension): also apply the node-feature Dropout that follows in the model, fused."""
        _require_cuda(x, type(self).__name__)
        # stand-alone use with int64 ids: out-of-vocabulary values are read as id 0 (memory-safe); the model path narrows
        # with an error flag instead (graph.batch_from_reference_tensors) and raises IndexError like torch's Embedding
        idx = x if x.dtype == torch.uint8 else ops.narrow_u8(x.contiguous(), self.dims)
        return Fn.embed_sum(self.table(), idx.contiguous(), self._offsets, dropout_p, seed)


class AtomEncoder(_FeatureEncoder):
    """``ogb...mol_encoder.AtomEncoder``: sum of 9 embedding lookups (models.py:175, 205)."""
    list_name, dims = "atom_embedding_list", ATOM_FEATURE_DIMS


class BondEncoder(_FeatureEncoder):
    """``ogb...mol_encoder.BondEncoder``: sum of 3 embedding lookups (models.py:176, 206)."""
    list_name, dims = "bond_embedding_list", BOND_FEATURE_DIMS

    def __init__(self, emb_dim: int):
        super().__init__(emb_dim)
        sel = torch.zeros(NUM_BOND_CODES, sum(BOND_FEATURE_DIMS))
        for code in range(NUM_BOND_CODES):
            a0, a1, a2 = code // 12, (code // 2) % 6, code % 2
            sel[code, a0] = 1.0
            sel[code, 5 + a1] = 1.0
            sel[code, 11 + a2] = 1.0
        self.register_buffer("_selector", sel, persistent=False)
        codes = torch.arange(NUM_BOND_CODES)
        self.register_buffer("_code_features", torch.stack([codes // 12, (codes // 2) % 6, codes % 2], 1), persistent=False)

    def table60(self) -> Tensor:
        """The 60-row table of every bond-feature combination with plain ATen ops (traceable: the custom-op face);
        the same sums in the same order as ``forward_table``'s selector product."""
        e = self.bond_embedding_list
        f = self._code_features
        return (e[0].weight[f[:, 0]] + e[1].weight[f[:, 1]]) + e[2].weight[f[:, 2]]

    def forward_table(self, codes: Tensor) -> EdgeTable:
        """Factored output: the 60-row table of every bond-feature combination + per-slot codes."""
        return EdgeTable(Fn.select_rows_matmul(self._selector, self.table()), codes)


# ------------------------------------------------------------------ batch norm / pooling
class BatchNorm(Module):
    """``torch_geometric.nn.BatchNorm`` (models.py:184): BatchNorm1d over all node rows, held as ``.module``."""

    def __init__(self, in_channels: int, eps: float = 1e-5, momentum: Optional[float] = 0.1, affine: bool = True,
                 track_running_stats: bool = True, allow_single_element: bool = False):
        super().__init__()
        if allow_single_element and not track_running_stats:
            raise ValueError("'allow_single_element' requires 'track_running_stats' to be set to `True`")
        self.module = BatchNorm1d(in_channels, eps, momentum, affine, track_running_stats)
        self.in_channels = in_channels
        self.allow_single_element = allow_single_element

    def reset_parameters(self):
        self.module.reset_parameters()

    def forward(self, x: Tensor, relu: bool = False, buckets=None, dropout_p: float = 0.0, seed: int = 0) -> Tensor:
        """``relu`` / ``dropout_p`` (gnnx extensions): fuse the ReLU that follows in the model (models.py:212) and the
        next layer's Dropout (models.py:209) into the same pass."""
        _require_cuda(x, "BatchNorm")
        if buckets is not None:
            return Fn.batch_norm(x, self.module, relu, buckets, dropout_p, seed)
        if self.allow_single_element and x.size(0) <= 1:
            was = self.module.training
            self.module.eval()
            try:
                return Fn.batch_norm(x, self.module, relu, None, dropout_p, seed)
            finally:
                self.module.train(was)
        return Fn.batch_norm(x, self.module, relu, None, dropout_p, seed)


class _SegmentAggregation(Module):
    mode = "add"

    def forward(self, x: Tensor, index: Optional[Tensor] = None, ptr: Optional[Tensor] = None,
                dim_size: Optional[int] = None, dim: int = -2) -> Tensor:
        """``aggr(x, batch)`` (models.py:219).  ``index`` must be sorted (PyG ``Batch.batch`` is); ``ptr`` may
        carry a prebuilt int32 ``graph_ptr`` to skip the host sync PyG pays for ``dim_size``."""
        _require_cuda(x, type(self).__name__)
        if x.dim() != 2 or dim not in (-2, 0):
            raise ValueError("gnnx aggregations reduce dim 0 of a 2-D tensor")
        if ptr is None:
            if index is None:
                raise ValueError("Aggregation requires 'index' to be specified")
            if dim_size is None:
                dim_size = int(index[-1].item()) + 1 if index.numel() > 0 else 0
            ptr = ops.graph_ptr_from_batch(index.to(torch.int64).contiguous(), dim_size)
        elif ptr.dtype != torch.int32:
            ptr = ptr.to(torch.int32)
        return Fn.segment_pool(x, ptr, self.mode)


class SumAggregation(_SegmentAggregation):
    mode = "add"


class MeanAggregation(_SegmentAggregation):
    mode = "mean"


class MaxAggregation(_SegmentAggregation):
    mode = "max"


# ------------------------------------------------------------------ PNA
class DegreeScalerAggregation(Module):
    """Holder of PNA's normalisers (PyG buffer names ``aggr_module.avg_deg_lin`` / ``avg_deg_log``)."""

    def __init__(self, aggr: Sequence[str], scaler: Sequence[str], deg: Tensor):
        super().__init__()
        if list(aggr) != _PNA_AGGREGATORS or list(scaler) != _PNA_SCALERS:
            raise NotImplementedError(
                f"gnnx PNAConv implements aggregators {_PNA_AGGREGATORS} with scalers {_PNA_SCALERS} "
                f"(gnnepcsaft/train/models.py:443-444); got {list(aggr)} / {list(scaler)}")
        deg = deg.to(torch.float)
        num_nodes = int(deg.sum())
        bin_degree = torch.arange(deg.numel())
        self.init_avg_deg_lin = float((bin_degree * deg).sum()) / num_nodes
        self.init_avg_deg_log = float(((bin_degree + 1).log() * deg).sum()) / num_nodes
        self.register_buffer("avg_deg_lin", torch.full((1,), self.init_avg_deg_lin))
        self.register_buffer("avg_deg_log", torch.full((1,), self.init_avg_deg_log))
        self._cache = None

    def avg_deg_log_value(self) -> float:
        from torch._subclasses.fake_tensor import is_fake
        if is_fake(self.avg_deg_log) or torch.compiler.is_compiling():
            # traced / exported: the buffer is a FakeTensor, its value cannot be read -- use the last value seen eagerly
            # (refreshed by __init__ and by every load_state_dict), which export bakes in as a constant
            return self._cache[1] if self._cache is not None else self.init_avg_deg_log
        key = (self.avg_deg_log.data_ptr(), self.avg_deg_log._version)
        if self._cache is None or self._cache[0] != key:
            if self._cache is not None and self.avg_deg_log.is_cuda and torch.cuda.is_current_stream_capturing():
                return self._cache[1]                                 # a host read would invalidate the capture
            self._cache = (key, float(self.avg_deg_log.item()))      # one sync per (re)load, not per step
        return self._cache[1]

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        self._cache = None
        if not self.avg_deg_log.is_meta:
            self.avg_deg_log_value()                                  # refresh the host copy while the value is readable


def _resolve_edges(edge_attr, edge_dim: int, g: GraphIndex, who: str):
    if isinstance(edge_attr, EdgeTable):
        return edge_attr.table, edge_attr.codes
    if edge_attr is None:
        raise NotImplementedError(f"{who}: the gnnx kernels implement the edge-feature variant (edge_dim is set in "
                                  "every reference config); edge_attr=None is not supported")
    if edge_attr.dim() != 2 or edge_attr.shape[0] != g.num_edges or edge_attr.shape[1] != edge_dim:
        raise ValueError(f"{who}: edge_attr must be [{g.num_edges}, {edge_dim}], got {tuple(edge_attr.shape)}")
    return edge_attr, None


class PNAConv(Module):
    """``torch_geometric.nn.PNAConv`` as built by ``get_conv`` (models.py:445-457)."""

    def __init__(self, in_channels: int, out_channels: int, aggregators: List[str], scalers: List[str], deg: Tensor,
                 edge_dim: Optional[int] = None, towers: int = 1, pre_layers: int = 1, post_layers: int = 1,
                 divide_input: bool = False, act: str = "relu", **kwargs):
        super().__init__()
        if not divide_input:
            raise NotImplementedError("gnnx PNAConv implements divide_input=True (gnnepcsaft/train/models.py:456)")
        if act != "relu":
            raise NotImplementedError("gnnx PNAConv implements act='relu'")
        if edge_dim is None:
            raise NotImplementedError("gnnx PNAConv implements the edge_dim variant")
        assert in_channels % towers == 0
        assert out_channels % towers == 0
        self.in_channels, self.out_channels = in_channels, out_channels
        self.aggregators, self.scalers = aggregators, scalers
        self.edge_dim, self.towers, self.divide_input = edge_dim, towers, divide_input
        self.pre_layers, self.post_layers = pre_layers, post_layers
        self.F_in = in_channels // towers
        self.F_out = out_channels // towers
        self.aggr_module = DegreeScalerAggregation(aggregators, scalers, deg)
        self.edge_encoder = Linear(edge_dim, self.F_in)
        self.pre_nns, self.post_nns = ModuleList(), ModuleList()
        for _ in range(towers):
            mods: List[Module] = [Linear(3 * self.F_in, self.F_in)]
            for _ in range(pre_layers - 1):
                mods += [ReLU(), Linear(self.F_in, self.F_in)]
            self.pre_nns.append(Sequential(*mods))
            mods = [Linear((len(aggregators) * len(scalers) + 1) * self.F_in, self.F_out)]
            for _ in range(post_layers - 1):
                mods += [ReLU(), Linear(self.F_out, self.F_out)]
            self.post_nns.append(Sequential(*mods))
        self.lin = Linear(out_channels, out_channels)

    def _flat_params(self) -> List[Tensor]:
        ps = [self.edge_encoder.weight, self.edge_encoder.bias]
        for nn in self.pre_nns:
            for m in nn:
                if isinstance(m, Linear):
                    ps += [m.weight, m.bias]
        for nn in self.post_nns:
            for m in nn:
                if isinstance(m, Linear):
                    ps += [m.weight, m.bias]
        return ps + [self.lin.weight, self.lin.bias]

    def _meta(self, g, codes, prep=None) -> PNAMeta:
        return PNAMeta(g, codes, self.towers, self.F_in, self.F_out, self.pre_layers, self.post_layers,
                       self.aggr_module.avg_deg_log_value(), prep)

    def prepare(self, table: Optional[Tensor], g: GraphIndex, n_groups: Optional[int] = None):
        """The parameter-only part of a call (``pna_fn.PNAPrep``), computed ahead of it -- GNNePCSAFT queues this for
        all its layers on the side stream.  ``table``: the factored bond-feature table, ``g``: the batch's index;
        ``n_groups``: the batch's degree-group count when the caller already has it (nothing of ``g`` is then touched:
        the call is pure parameter arithmetic and can be captured into a CUDA graph)."""
        meta = self._meta(g, None)
        if n_groups is None:
            bk = g.degree_buckets() if grouped_ok(meta) else None
            n_groups = bk.n_groups if bk is not None else 0
        elif not grouped_ok(meta):
            n_groups = 0
        with torch.no_grad():
            return pna_prepare(meta, self._flat_params(), table, n_groups)

    def bucket_rows(self, g: GraphIndex):
        """The batch's degree buckets when this layer takes the degree-grouped path on ``g`` (else ``None``)."""
        return g.degree_buckets() if grouped_ok(self._meta(g, None)) else None

    def forward(self, x: Tensor, edge_index: Optional[Tensor], edge_attr: Union[Tensor, EdgeTable, None] = None,
                index: Optional[GraphIndex] = None, prep=None, bucket_out: bool = False) -> Tensor:
        """``bucket_out`` (only with ``bucket_rows(g)`` not ``None``): leave the result in degree-bucket order
        [capacity, out_channels] for a consumer that reads rows through the bucket map (``BatchNorm(..., buckets=)``)."""
        _require_cuda(x, "PNAConv")
        if x.dim() != 2 or x.shape[1] != self.in_channels:
            raise ValueError(f"PNAConv: x must be [N, {self.in_channels}], got {tuple(x.shape)}")
        g = index if index is not None else graph_index_for(edge_index, x.shape[0])
        efeat, codes = _resolve_edges(edge_attr, self.edge_dim, g, "PNAConv")
        meta = self._meta(g, codes, prep)
        meta.bucket_out = bucket_out
        return PNAConvFn.apply(x, efeat, meta, *self._flat_params())


# ------------------------------------------------------------------ GATv2
class GATv2Conv(Module):
    """``torch_geometric.nn.GATv2Conv`` as built by ``get_conv`` (models.py:487-495)."""

    def __init__(self, in_channels: int, out_channels: int, heads: int = 1, concat: bool = True,
                 negative_slope: float = 0.2, dropout: float = 0.0, add_self_loops: bool = True,
                 edge_dim: Optional[int] = None, fill_value: Union[float, str] = "mean", bias: bool = True,
                 share_weights: bool = False, residual: bool = False, **kwargs):
        super().__init__()
        if not concat or share_weights or residual or fill_value != "mean" or edge_dim is None:
            raise NotImplementedError("gnnx GATv2Conv implements concat=True, share_weights=False, residual=False, "
                                      "fill_value='mean' with edge features (gnnepcsaft/train/models.py:487-495)")
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.concat, self.negative_slope, self.dropout = concat, negative_slope, dropout
        self.add_self_loops, self.edge_dim, self.fill_value = add_self_loops, edge_dim, fill_value
        self.lin_l = Linear(in_channels, heads * out_channels, bias=bias)
        self.lin_r = Linear(in_channels, heads * out_channels, bias=bias)
        self.att = Parameter(torch.empty(1, heads, out_channels))
        self.lin_edge = Linear(edge_dim, heads * out_channels, bias=False)
        self.bias = Parameter(torch.empty(heads * out_channels)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        for lin in (self.lin_l, self.lin_r, self.lin_edge):
            torch.nn.init.xavier_uniform_(lin.weight)
            if lin.bias is not None:
                torch.nn.init.zeros_(lin.bias)
        torch.nn.init.xavier_uniform_(self.att)
        if self.bias is not None:
            torch.nn.init.zeros_(self.bias)

    def forward(self, x: Tensor, edge_index: Optional[Tensor], edge_attr: Union[Tensor, EdgeTable, None] = None,
                index: Optional[GraphIndex] = None) -> Tensor:
        _require_cuda(x, "GATv2Conv")
        if x.dim() != 2 or x.shape[1] != self.in_channels:
            raise ValueError(f"GATv2Conv: x must be [N, {self.in_channels}], got {tuple(x.shape)}")
        g = index if index is not None else graph_index_for(edge_index, x.shape[0])
        efeat, codes = _resolve_edges(edge_attr, self.edge_dim, g, "GATv2Conv")
        p = self.dropout if self.training else 0.0           # F.dropout(alpha, p, training) inside the kernel
        meta = GATMeta(g, codes, self.heads, self.out_channels, self.negative_slope, self.add_self_loops, p,
                       ops.dropout_seed() if p > 0.0 else 0)
        return GATv2ConvFn.apply(x, efeat, meta, self.lin_l.weight, self.lin_l.bias, self.lin_r.weight,
                                 self.lin_r.bias, self.lin_edge.weight, self.att, self.bias)


# ------------------------------------------------------------------ TransformerConv
class TransformerConv(Module):
    """``torch_geometric.nn.TransformerConv`` as built by ``get_conv`` (models.py:497-510; the third conv of
    configs/search_space.py:15): ``concat=True, beta=False, root_weight=True, bias=True`` with edge features."""

    def __init__(self, in_channels: int, out_channels: int, heads: int = 1, concat: bool = True, beta: bool = False,
                 dropout: float = 0.0, edge_dim: Optional[int] = None, bias: bool = True, root_weight: bool = True,
                 **kwargs):
        super().__init__()
        if not concat or beta or not root_weight or not bias or edge_dim is None:
            raise NotImplementedError("gnnx TransformerConv implements concat=True, beta=False, root_weight=True, "
                                      "bias=True with edge features (gnnepcsaft/train/models.py:497-510)")
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.concat, self.beta, self.root_weight = concat, beta, root_weight
        self.dropout, self.edge_dim = dropout, edge_dim
        self.lin_key = Linear(in_channels, heads * out_channels)
        self.lin_query = Linear(in_channels, heads * out_channels)
        self.lin_value = Linear(in_channels, heads * out_channels)
        self.lin_edge = Linear(edge_dim, heads * out_channels, bias=False)
        self.lin_skip = Linear(in_channels, heads * out_channels, bias=bias)

    def reset_parameters(self):
        for lin in (self.lin_key, self.lin_query, self.lin_value, self.lin_edge, self.lin_skip):
            lin.reset_parameters()

    def forward(self, x: Tensor, edge_index: Optional[Tensor], edge_attr: Union[Tensor, EdgeTable, None] = None,
                index: Optional[GraphIndex] = None) -> Tensor:
        _require_cuda(x, "TransformerConv")
        if x.dim() != 2 or x.shape[1] != self.in_channels:
            raise ValueError(f"TransformerConv: x must be [N, {self.in_channels}], got {tuple(x.shape)}")
        g = index if index is not None else graph_index_for(edge_index, x.shape[0])
        efeat, codes = _resolve_edges(edge_attr, self.edge_dim, g, "TransformerConv")
        p = self.dropout if self.training else 0.0           # F.dropout(alpha, p, training) inside the kernel
        meta = TConvMeta(g, codes, self.heads, self.out_channels, p, ops.dropout_seed() if p > 0.0 else 0)
        return TransformerConvFn.apply(x, efeat, meta, self.lin_query.weight, self.lin_query.bias, self.lin_key.weight,
                                       self.lin_key.bias, self.lin_value.weight, self.lin_value.bias,
                                       self.lin_edge.weight, self.lin_skip.weight, self.lin_skip.bias)
