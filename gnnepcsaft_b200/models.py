"""Drop-in ``GNNePCSAFT`` (reference: gnnepcsaft/train/models.py:159-254, 441-606).

Same config keys, submodule / state-dict names, ``forward(x, edge_index, edge_attr, batch)`` and
``pred_with_bounds(data)`` as the reference model; the body runs on the gnnx sm_100a kernels:
the CSR is built once per batch, bond features are kept factored (60-row table + codes), and
BatchNorm + ReLU are one kernel per layer.
"""

from __future__ import annotations

import math
from typing import Any, Optional, Union

import torch
import torch.nn.functional as F
from torch import Tensor
from torch.nn import BatchNorm1d, Dropout, Linear, Module, ModuleList, ReLU, Sequential

from ._lib import lib
from . import functional as Fn
from . import nn as gnn
from . import ops, streams
from .graph import Batch, batch_from_reference_tensors


# Debug tap: when set to a list, the MLP head appends its two BatchNorm+ReLU outputs (tests teacher-force their sign
# pattern onto the oracle's head ReLUs).
CAPTURE_HEAD = None
# Debug tap: when set to a list, every training-mode forward appends (p, [one seed per conv layer]) of its node-feature
# dropout draws (tests rebuild the keep-masks with ops.dropout_keep_mask and replay them on the oracle).
NODE_DROPOUT_SEEDS = None


_FORCE_CUSTOM_OPS = False


def use_custom_ops(flag: bool) -> None:
    """Route ``GNNePCSAFT.forward`` through the ``torch.ops.gnnx.*`` custom-op face (``torch_ops.py``) in eager mode too.
    Tracing (``torch.export``, FakeTensor inputs) always takes it."""
    global _FORCE_CUSTOM_OPS
    _FORCE_CUSTOM_OPS = bool(flag)


def _traced(x: Tensor) -> bool:
    from torch._subclasses.fake_tensor import is_fake
    return _FORCE_CUSTOM_OPS or is_fake(x) or torch.compiler.is_compiling() or torch.compiler.is_exporting()


def get_conv(config: dict) -> Module:
    """Conv factory, reference models.py:441-584 (the convs any file in ``configs/`` selects)."""
    aggregators = ["mean", "min", "max", "std"]
    scalers = ["identity", "amplification", "attenuation"]
    if config["conv"] == "PNA":
        return gnn.PNAConv(
            in_channels=config["hidden_dim"], out_channels=config["hidden_dim"], aggregators=aggregators,
            scalers=scalers, deg=torch.tensor(config["deg"], dtype=torch.long), edge_dim=config["hidden_dim"],
            towers=config["towers"], pre_layers=config["pre_layers"], post_layers=config["post_layers"],
            divide_input=True)
    if config["conv"] == "GATv2":
        assert config["hidden_dim"] % config["heads"] == 0, "hidden_dim must be divisible by heads"
        return gnn.GATv2Conv(
            in_channels=config["hidden_dim"], out_channels=config["hidden_dim"] // config["heads"],
            heads=config["heads"], concat=True, dropout=config["dropout"], edge_dim=config["hidden_dim"],
            add_self_loops=config["add_self_loops"])
    if config["conv"] == "Transformer":
        assert config["hidden_dim"] % config["heads"] == 0, "hidden_dim must be divisible by heads"
        return gnn.TransformerConv(
            in_channels=config["hidden_dim"], out_channels=config["hidden_dim"] // config["heads"],
            heads=config["heads"], concat=True, dropout=config["dropout"], edge_dim=config["hidden_dim"])
    raise ValueError(f"Unsupported convolution: {config['conv']}.")


def get_global_pool(config: dict) -> Module:
    """Pool factory, reference models.py:587-595."""
    if config["global_pool"] == "mean":
        return gnn.MeanAggregation()
    if config["global_pool"] == "max":
        return gnn.MaxAggregation()
    if config["global_pool"] == "add":
        return gnn.SumAggregation()
    raise ValueError(f"Unsupported global pooling: {config['global_pool']}.")


class GNNePCSAFT(Module):
    """Graph neural network to predict PC-SAFT parameters (reference models.py:159-254)."""

    def __init__(self, config: dict):
        super().__init__()
        self.convs = ModuleList()
        self.batch_norms = ModuleList()
        self.lower_bounds = torch.tensor([1.0, 1.9, 50.0, -1 * math.log10(0.9), math.log10(200.0)])
        self.upper_bounds = torch.tensor([25.0, 4.5, 550.0, -1 * math.log10(0.0001), math.log10(5000.0)])
        self.num_para = config["num_para"]
        self.node_embed = gnn.AtomEncoder(config["hidden_dim"])
        self.edge_embed = gnn.BondEncoder(config["hidden_dim"])
        self.dropout = Dropout(p=config["dropout"])
        self.global_pool = get_global_pool(config)
        self.global_pool_type = config["global_pool"]
        for _ in range(config["propagation_depth"]):
            self.convs.append(get_conv(config))
            self.batch_norms.append(gnn.BatchNorm(config["hidden_dim"]))
        h = config["hidden_dim"]
        self.mlp = Sequential(
            Linear(h, h // 2), BatchNorm1d(h // 2), ReLU(),
            Linear(h // 2, h // 4), BatchNorm1d(h // 4), ReLU(),
            Linear(h // 4, config["num_para"]),
        )
        self._bounds_dev = None
        self._prep_cache = None
        self._prep_graph = None
        # PNA: the dataset's largest in-degree (the length of the `deg` histogram the layers were built from) bounds every
        # batch of that dataset, so the reference-signature forward can bucket nodes by degree without reading the batch's
        # maximum back (one host sync per batch less).  A batch that breaks the bound is flagged on the device and raises
        # at the next forward / validate(); None switches the hint off.
        deg = config.get("deg") if config.get("conv") == "PNA" else None
        self.degree_hint = (len(deg) - 1) if deg is not None and len(deg) >= 1 else None
        self._hinted = None

    # -- reference signature -------------------------------------------------------------
    def forward(self, x: Tensor, edge_index: Tensor, edge_attr: Tensor, batch: Union[Tensor, None],
                num_graphs: Optional[int] = None) -> Tensor:
        """Forward pass of the model (same inputs / output as the reference's ``forward``)."""
        if not x.is_cuda:
            raise RuntimeError("GNNePCSAFT (gnnx): inputs must be CUDA tensors; there is no CPU fallback")
        if _traced(x):
            return self.forward_custom_ops(x, edge_index, edge_attr, batch, num_graphs)
        b = batch_from_reference_tensors(x.contiguous(), edge_index.contiguous(), edge_attr.contiguous(), batch,
                                         num_graphs)
        self.check_pending()
        if self.degree_hint is not None and b.index.max_degree is None:
            if b.index._buckets is None:                 # (a caller may have bucketed ahead, with the same hint)
                b.index.degree_hint = self.degree_hint
            self._hinted = b.index
        return self.forward_batch(b)

    def check_pending(self) -> None:
        """Raise if the PREVIOUS reference-signature batch broke the degree hint (its status words are read now: that
        batch's index kernels finished long ago, so this does not stall the launch queue)."""
        g, self._hinted = self._hinted, None
        if g is not None:
            g.read_status_async()

    # -- the torch.library face: every step is a torch.ops.gnnx.* custom op or a plain ATen op ----------------
    def forward_custom_ops(self, x: Tensor, edge_index: Tensor, edge_attr: Tensor, batch: Union[Tensor, None],
                           num_graphs: Optional[int] = None) -> Tensor:
        """Same result as ``forward`` through ``torch.ops.gnnx.*`` only (``torch_ops.py``): no host read, no Python
        object crosses an op boundary, shapes may be symbolic -- what ``torch.export.export(model, (x, edge_index,
        edge_attr, None), dynamic_shapes=...)`` (reference demo/utils.py:890-915) traces.  Differentiable."""
        from . import torch_ops  # noqa: F401  (registers the ops)
        T = torch.ops.gnnx
        N = x.shape[0]
        index = T.csr_build(edge_index, N)
        status = index[7]
        x8 = T.narrow_u8(x, list(gnn.ATOM_FEATURE_DIMS), status, ops.ST_ATOM)
        ea8 = T.narrow_u8(edge_attr, list(gnn.BOND_FEATURE_DIMS), status, ops.ST_BOND)
        codes = T.edge_codes(ea8, index[2], status)
        table = self.edge_embed.table60()
        p = self.dropout.p if self.training else 0.0
        L = len(self.convs)
        seeds = [ops.dropout_seed() if p > 0.0 else 0 for _ in range(L)]
        h = T.embed_sum(self.node_embed.table(), x8, self.node_embed._offsets, p, seeds[0] if L else 0)
        for l, (conv, bn) in enumerate(zip(self.convs, self.batch_norms)):
            pn, sn = (p, seeds[l + 1]) if l + 1 < L else (0.0, 0)
            if isinstance(conv, gnn.PNAConv):
                h = T.pna_conv(h, table, codes, index, conv._flat_params(), conv.towers, conv.pre_layers, conv.post_layers,
                               conv.aggr_module.avg_deg_log_value(), -1)
            elif isinstance(conv, gnn.GATv2Conv):
                pa = conv.dropout if self.training else 0.0
                h = T.gatv2_conv(h, table, codes, index, [conv.lin_l.weight, conv.lin_l.bias, conv.lin_r.weight, conv.lin_r.bias,
                                                          conv.lin_edge.weight, conv.att, conv.bias], conv.heads,
                                 conv.out_channels, conv.negative_slope, conv.add_self_loops, pa,
                                 ops.dropout_seed() if pa > 0.0 else 0)
            else:
                pa = conv.dropout if self.training else 0.0
                h = T.transformer_conv(h, table, codes, index, [conv.lin_query.weight, conv.lin_query.bias, conv.lin_key.weight,
                                                                conv.lin_key.bias, conv.lin_value.weight, conv.lin_value.bias,
                                                                conv.lin_edge.weight, conv.lin_skip.weight, conv.lin_skip.bias],
                                       conv.heads, conv.out_channels, pa, ops.dropout_seed() if pa > 0.0 else 0)
            h = self._bn_op(h, bn.module, pn, sn)
        if batch is not None:
            if num_graphs is None:
                raise ValueError("GNNePCSAFT (custom-op face): pass num_graphs= together with a batch vector (reading "
                                 "int(batch[-1]) + 1 back is a host sync a traced graph cannot contain)")
            ptr = T.graph_ptr(batch, num_graphs)
        else:
            ptr = (torch.arange(2, device=x.device) * N).to(torch.int32)          # [0, N]: one graph
        h = T.segment_pool(h, ptr, ops.POOL_MODES[self.global_pool_type])
        m = self.mlp
        h = self._bn_op(T.linear(h, m[0].weight, m[0].bias, False), m[1], 0.0, 0)
        h = self._bn_op(T.linear(h, m[3].weight, m[3].bias, False), m[4], 0.0, 0)
        return T.linear(h, m[6].weight, m[6].bias, False)

    @staticmethod
    def _bn_op(h: Tensor, bn: BatchNorm1d, dropout_p: float, seed: int) -> Tensor:
        training = bn.training or not bn.track_running_stats
        if training and bn.track_running_stats and bn.num_batches_tracked is not None:
            bn.num_batches_tracked.add_(1)
        momentum = 0.1 if bn.momentum is None else bn.momentum
        out = torch.ops.gnnx.bn_relu(h, bn.weight, bn.bias, bn.running_mean, bn.running_var, training, momentum, bn.eps,
                                     True, dropout_p, seed)
        if training and bn.track_running_stats:
            with torch.no_grad():
                bn.running_mean.copy_(out[3])
                bn.running_var.copy_(out[4])
        return out[0]

    # -- fast path: everything prebuilt on the device --------------------------------------
    def forward_batch(self, b: Batch) -> Tensor:
        g = b.index
        st = g.err_flag
        x_idx = b.x if b.x.dtype == torch.uint8 else ops.narrow_u8(b.x, gnn.ATOM_FEATURE_DIMS, st[ops.ST_ATOM:ops.ST_ATOM + 1])
        ea_u8 = b.edge_attr if b.edge_attr.dtype == torch.uint8 else ops.narrow_u8(
            b.edge_attr, gnn.BOND_FEATURE_DIMS, st[ops.ST_BOND:ops.ST_BOND + 1])
        edges = self.edge_embed.forward_table(ops.edge_codes(ea_u8.contiguous(), g.csr_eid, st[ops.ST_BOND:ops.ST_BOND + 1]))
        if self.training and torch.is_grad_enabled():
            ops.code_order_ahead(edges.codes, edges.table.shape[0])     # the backward's slot ordering, under the forward
        preps = self._prepare_convs(edges, g)
        # Dropout(p) in front of every conv (reference models.py:209) is fused into the kernel that PRODUCES that conv's
        # input: the embedding sum for layer 0, the BatchNorm+ReLU apply of layer l-1 for layer l.  No mask tensor: the
        # draw is a function of (seed, element index) that the backward replays.  NODE_DROPOUT_SEEDS taps the seeds.
        p = self.dropout.p if self.training else 0.0
        L = len(self.convs)
        seeds = [ops.dropout_seed() if p > 0.0 else 0 for _ in range(L)]
        if NODE_DROPOUT_SEEDS is not None:
            NODE_DROPOUT_SEEDS.append((p, list(seeds)))
        x = self.node_embed(x_idx, dropout_p=p, seed=seeds[0]) if L else self.node_embed(x_idx)
        for l, (conv, batch_norm, prep) in enumerate(zip(self.convs, self.batch_norms, preps)):
            pn, sn = (p, seeds[l + 1]) if l + 1 < L else (0.0, 0)     # the next conv's dropout; none after the last layer
            if isinstance(conv, gnn.PNAConv):
                # degree-grouped layers hand their rows to the batch norm in bucket order (no extra pass back to node order)
                bk = conv.bucket_rows(g)
                out = conv(x, b.edge_index, edges, index=g, prep=prep, bucket_out=bk is not None)
                x = batch_norm(out, relu=True, buckets=bk, dropout_p=pn, seed=sn)
            else:
                x = batch_norm(conv(x, b.edge_index, edges, index=g), relu=True, dropout_p=pn, seed=sn)
        if b.batch is not None or b.graph_ptr is not None:
            x = self.global_pool(x, ptr=b.graph_ptr)
        else:
            n = x.shape[0]
            ptr = torch.tensor([0, n], dtype=torch.int32, device=x.device)
            x = Fn.segment_pool(x, ptr, self.global_pool_type)
        return self._head(x)

    def _prepare_convs(self, edges, g):
        """Queue the parameter-only work of every PNAConv layer (fused / packed weights, the 60-row edge-term tables,
        the per-degree weights) on the side stream, ahead of the layer loop: ~5 tiny kernels per layer that would
        otherwise each cost a launch latency between the HBM-bound kernels.  Each layer waits on its own event.
        Without gradients and outside training the result is kept until a parameter changes (bulk screening runs
        thousands of batches on fixed weights)."""
        L = len(self.convs)
        sd = streams.side(edges.table.device)
        if sd is None or not all(isinstance(c, gnn.PNAConv) for c in self.convs):
            return [None] * L
        reuse = not torch.is_grad_enabled() and not self.training
        if not reuse and not streams.PREP_AHEAD:
            return [None] * L
        if not reuse and streams.PREP_GRAPH:
            return self._prepare_convs_graph(edges, g, sd)
        key = None
        if reuse:
            bk = g.degree_buckets()
            key = (ops.param_epoch(), bk.n_groups if bk is not None else 0) + tuple(
                (p.data_ptr(), p._version) for m in (self.convs, self.edge_embed) for p in m.parameters())
            if self._prep_cache is not None and self._prep_cache[0] == key:
                return self._prep_cache[1]
        sd.begin()
        try:
            preps = []
            for conv in self.convs:
                p = conv.prepare(edges.table.detach(), g)
                p.event = sd.mark()
                preps.append(p)
        finally:
            sd.end()
        self._prep_cache = (key, preps) if reuse else None
        return preps

    def _prepare_convs_graph(self, edges, g, sd):
        """Training: the same parameter-only work (~10 tiny kernels per layer: packing, 60-row tables, W_lin W_last, the
        per-degree weights) replayed from ONE CUDA graph on the side stream.  Everything it reads is static -- the
        parameters' storage and a persistent copy of the bond table -- and so is what it writes (the graph's pool), so
        a step costs one copy and one graph launch instead of ~60 launches; the layers wait on one event.  Eager on
        the first step with a given key (warm-up; also what allocates the kernels' caches), captured on the second.
        Inside somebody else's capture (a whole step being captured) the layers fall back to inline preparation."""
        L = len(self.convs)
        if torch.cuda.is_current_stream_capturing():
            return [None] * L
        bk = g.degree_buckets()
        n_groups = bk.n_groups if bk is not None else 0
        table = edges.table.detach()
        key = (n_groups, tuple(table.shape)) + tuple(
            p.data_ptr() for m in (self.convs, self.edge_embed) for p in m.parameters())
        ent = self._prep_graph
        if ent is None or ent["key"] != key:
            ent = self._prep_graph = dict(key=key, graph=None, table=torch.empty_like(table), preps=None, launches=0)
            sd.begin()                                  # first step with this key: eager, on the side stream
            try:
                preps = [conv.prepare(table, g, n_groups=n_groups) for conv in self.convs]
                ev = sd.mark()
            finally:
                sd.end()
            for p in preps:
                p.event = ev
            return preps
        if ent["graph"] is None:
            ent["table"].copy_(table)
            graph = torch.cuda.CUDAGraph()
            n0 = lib().launch_count()
            # thread_local: a loader thread allocating pinned memory meanwhile must not invalidate the capture
            with torch.cuda.graph(graph, stream=sd.stream, capture_error_mode="thread_local"):
                ent["preps"] = [conv.prepare(ent["table"], g, n_groups=n_groups) for conv in self.convs]
            ent["graph"] = graph
            ent["launches"] = lib().launch_count() - n0
        sd.begin()
        try:
            ent["table"].copy_(table)
            ent["graph"].replay()
            lib().graph_replayed += ent["launches"]
            ev = sd.mark()
        finally:
            sd.end()
        for p in ent["preps"]:
            p.event = ev
        return ent["preps"]

    def _head(self, x: Tensor) -> Tensor:
        m = self.mlp
        x = Fn.batch_norm(Fn.linear(x, m[0].weight, m[0].bias), m[1], relu=True)
        if CAPTURE_HEAD is not None:
            CAPTURE_HEAD.append(x.detach())
        x = Fn.batch_norm(Fn.linear(x, m[3].weight, m[3].bias), m[4], relu=True)
        if CAPTURE_HEAD is not None:
            CAPTURE_HEAD.append(x.detach())
        return Fn.linear(x, m[6].weight, m[6].bias)

    def pred_with_bounds(self, data: Any) -> Tensor:
        """Forward pass of the model with bounds (reference models.py:229-254)."""
        if isinstance(data, Batch) and data.index is not None:
            params = self.forward_batch(data)
            dev = params.device
        else:
            x, edge_index, edge_attr, batch = data.x, data.edge_index, data.edge_attr, getattr(data, "batch", None)
            if not (isinstance(x, Tensor) and isinstance(edge_index, Tensor) and isinstance(edge_attr, Tensor)):
                raise ValueError("Invalid input data")
            params = self.forward(x, edge_index, edge_attr, batch, getattr(data, "num_graphs", None) or None)
            dev = x.device
        if self._bounds_dev is None or self._bounds_dev[0].device != dev:
            sl = slice(0, 3) if self.num_para == 3 else slice(3, None)
            self._bounds_dev = (self.lower_bounds[sl].to(dev), self.upper_bounds[sl].to(dev))
        return params.clip(self._bounds_dev[0], self._bounds_dev[1])


def create_model(config: dict, deg: list) -> GNNePCSAFT:
    """Reference ``create_model`` (models.py:598-606) without the Lightning wrapper."""
    config["deg"] = deg
    if config.get("model", "gnn").lower() == "gnn":
        return GNNePCSAFT(config)
    raise ValueError(f"Unsupported model: {config['model']}.")
