"""``torch.library`` custom-op face of the C ABI: ``torch.ops.gnnx.*``.

SURVEY 8(b) "Python side": every op is a ``torch.library.custom_op`` whose body calls the C entry points of
``include/gnnx.h`` (through ``ops.py``), with

  * ``register_fake``      -- shape / dtype inference on FakeTensors with symbolic sizes, so that
                              ``torch.export.export(model, (x, edge_index, edge_attr, None), dynamic_shapes=...)`` -- the call the
                              reference makes in ``demo/utils.py:890-915`` -- traces the model without touching the GPU;
  * ``register_autograd``  -- a backward formula for every differentiable op.

Two layers of ops:

  kernel-level   ``csr_build``, ``narrow_u8``, ``edge_codes``, ``embed_sum``, ``linear``, ``bn_relu``,
                 ``segment_pool`` (+ their ``*_bwd`` twins, themselves custom ops so a backward can be traced too);
  layer-level    ``pna_conv``, ``gatv2_conv``, ``transformer_conv``: one op per conv layer, taking the layer's
                 parameters as ``Tensor[]`` and the batch index as ``Tensor[]`` (what ``csr_build`` returns).

The eager TRAINING path of ``GNNePCSAFT`` does not go through this file: it calls the same C entry points from the fused
``autograd.Function`` nodes (``pna_fn`` / ``gat_fn`` / ``tconv_fn``), which keep their activations between forward and
backward and cost one Python dispatch per layer; a ``torch.library`` dispatch costs ~20 us, which at ~450 entry-point
calls per cfg2 step would make the step host-bound.  The layer-level custom ops are differentiable all the same: their
``register_autograd`` backward re-runs the fused node under ``enable_grad`` (activation recomputation, one extra
forward) -- the right trade for a traced / exported graph, where nothing may be stashed outside the graph.
``GNNePCSAFT.forward`` switches to these ops by itself whenever it is being traced (FakeTensor inputs /
``torch.compiler.is_exporting()``), and ``gnnepcsaft_b200.models.use_custom_ops(True)`` forces them in eager mode
(tests compare the two faces).
"""

from __future__ import annotations

from typing import List, Optional, Tuple

import torch
from torch import Tensor

from . import ops

custom_op = torch.library.custom_op
INDEX_PARTS = 8            # rowptr, csr_src, csr_eid, pos_of_eid, colptr, csc_pos, csc_dst, status


# ------------------------------------------------------------------ collator
@custom_op("gnnx::csr_build", mutates_args=())
def csr_build(edge_index: Tensor, num_nodes: int) -> List[Tensor]:
    return list(ops.csr_build(edge_index.contiguous(), int(num_nodes)))


@csr_build.register_fake
def _(edge_index, num_nodes):
    E = edge_index.shape[1]
    i32 = dict(dtype=torch.int32, device=edge_index.device)
    n1 = num_nodes + 1
    return [edge_index.new_empty((n1,), **i32), edge_index.new_empty((E,), **i32), edge_index.new_empty((E,), **i32),
            edge_index.new_empty((E,), **i32), edge_index.new_empty((n1,), **i32), edge_index.new_empty((E,), **i32),
            edge_index.new_empty((E,), **i32), edge_index.new_empty((ops.STATUS_WORDS,), **i32)]


@custom_op("gnnx::narrow_u8", mutates_args=("status",))
def narrow_u8(t: Tensor, limits: List[int], status: Tensor, slot: int) -> Tensor:
    return ops.narrow_u8(t.contiguous(), list(limits), status[slot:slot + 1])


@narrow_u8.register_fake
def _(t, limits, status, slot):
    return t.new_empty(t.shape, dtype=torch.uint8)


@custom_op("gnnx::edge_codes", mutates_args=("status",))
def edge_codes(edge_attr_u8: Tensor, csr_eid: Tensor, status: Tensor) -> Tensor:
    return ops.edge_codes(edge_attr_u8.contiguous(), csr_eid, status[ops.ST_BOND:ops.ST_BOND + 1])


@edge_codes.register_fake
def _(edge_attr_u8, csr_eid, status):
    return csr_eid.new_empty(csr_eid.shape, dtype=torch.uint8)


@custom_op("gnnx::graph_ptr", mutates_args=())
def graph_ptr(batch: Tensor, num_graphs: int) -> Tensor:
    return ops.graph_ptr_from_batch(batch.contiguous(), int(num_graphs))


@graph_ptr.register_fake
def _(batch, num_graphs):
    return batch.new_empty((num_graphs + 1,), dtype=torch.int32)


# ------------------------------------------------------------------ embedding sums
@custom_op("gnnx::embed_sum", mutates_args=())
def embed_sum(table: Tensor, idx_u8: Tensor, offsets: Tensor, dropout_p: float, seed: int) -> Tensor:
    return ops.embed_sum_fwd(idx_u8.contiguous(), offsets, table.contiguous(), dropout_p, seed)


@embed_sum.register_fake
def _(table, idx_u8, offsets, dropout_p, seed):
    return table.new_empty((idx_u8.shape[0], table.shape[1]))


@custom_op("gnnx::embed_sum_bwd", mutates_args=())
def embed_sum_bwd(dout: Tensor, idx_u8: Tensor, offsets: Tensor, table_rows: int, dropout_p: float, seed: int) -> Tensor:
    return ops.embed_sum_bwd(idx_u8, offsets, int(table_rows), dout.contiguous(), dropout_p, seed)


@embed_sum_bwd.register_fake
def _(dout, idx_u8, offsets, table_rows, dropout_p, seed):
    return dout.new_empty((table_rows, dout.shape[1]))


def _embed_setup(ctx, inputs, output):
    table, idx_u8, offsets, dropout_p, seed = inputs
    ctx.save_for_backward(idx_u8, offsets)
    ctx.rows, ctx.drop = table.shape[0], (dropout_p, seed)


def _embed_bwd(ctx, dout):
    idx_u8, offsets = ctx.saved_tensors
    return torch.ops.gnnx.embed_sum_bwd(dout, idx_u8, offsets, ctx.rows, ctx.drop[0], ctx.drop[1]), None, None, None, None


embed_sum.register_autograd(_embed_bwd, setup_context=_embed_setup)


# ------------------------------------------------------------------ Linear (+ ReLU)
@custom_op("gnnx::linear", mutates_args=())
def linear(x: Tensor, weight: Tensor, bias: Optional[Tensor], relu: bool) -> Tensor:
    return ops.gemm(x.contiguous(), weight, trans_b=True, bias=bias, relu=relu)


@linear.register_fake
def _(x, weight, bias, relu):
    return x.new_empty((x.shape[0], weight.shape[0]))


@custom_op("gnnx::linear_bwd", mutates_args=())
def linear_bwd(dy: Tensor, x: Tensor, weight: Tensor, y: Optional[Tensor]) -> List[Tensor]:
    """-> [dx, dweight, dbias]; ``y``: the ReLU'd output (its backward is applied to dy first) or None."""
    dy = dy.contiguous()
    if y is not None:
        dy = ops.relu_bwd(dy, y)
    return [ops.gemm(dy, weight), ops.gemm(dy, x, trans_a=True), ops.colsum(dy)]


@linear_bwd.register_fake
def _(dy, x, weight, y):
    return [x.new_empty(x.shape), weight.new_empty(weight.shape), weight.new_empty((weight.shape[0],))]


def _linear_setup(ctx, inputs, output):
    x, weight, bias, relu = inputs
    ctx.has_bias = bias is not None
    ctx.save_for_backward(x, weight, output if relu else None)


def _linear_bwd(ctx, dy):
    x, weight, y = ctx.saved_tensors
    dx, dw, db = torch.ops.gnnx.linear_bwd(dy, x, weight, y)
    return dx, dw, (db if ctx.has_bias else None), None


linear.register_autograd(_linear_bwd, setup_context=_linear_setup)


# ------------------------------------------------------------------ BatchNorm + ReLU (+ the next layer's Dropout)
@custom_op("gnnx::bn_relu", mutates_args=())
def bn_relu(x: Tensor, weight: Optional[Tensor], bias: Optional[Tensor], running_mean: Optional[Tensor],
            running_var: Optional[Tensor], training: bool, momentum: float, eps: float, relu: bool, dropout_p: float,
            seed: int) -> List[Tensor]:
    """-> [y, save_mean, save_rstd, new_running_mean, new_running_var].  Functional (an op with an autograd formula may
    not mutate its inputs): the running statistics come back as NEW tensors which the caller copies into the module's
    buffers; without running statistics the last two are empty."""
    rm = running_mean.clone() if running_mean is not None else None
    rv = running_var.clone() if running_var is not None else None
    y, mean, rstd = ops.bn_relu_fwd(x.contiguous(), weight, bias, rm, rv, training, momentum, eps, relu, None, dropout_p, seed)
    empty = x.new_empty((0,))
    return [y, mean, rstd, rm if rm is not None else empty, rv if rv is not None else empty.clone()]


@bn_relu.register_fake
def _(x, weight, bias, running_mean, running_var, training, momentum, eps, relu, dropout_p, seed):
    C = x.shape[1]
    return [x.new_empty(x.shape), x.new_empty((C,)), x.new_empty((C,)),
            x.new_empty((C,) if running_mean is not None else (0,)), x.new_empty((C,) if running_var is not None else (0,))]


@custom_op("gnnx::bn_relu_bwd", mutates_args=())
def bn_relu_bwd(dy: Tensor, x: Tensor, weight: Optional[Tensor], bias: Optional[Tensor], mean: Tensor, rstd: Tensor,
                relu: bool, training: bool, dropout_p: float, seed: int) -> List[Tensor]:
    return list(ops.bn_relu_bwd(dy.contiguous(), x, None, weight, mean, rstd, relu, None, bias=bias, training=training,
                                dropout_p=dropout_p, seed=seed))


@bn_relu_bwd.register_fake
def _(dy, x, weight, bias, mean, rstd, relu, training, dropout_p, seed):
    return [x.new_empty(x.shape), x.new_empty((x.shape[1],)), x.new_empty((x.shape[1],))]


def _bn_setup(ctx, inputs, output):
    x, weight, bias, _rm, _rv, training, _mom, _eps, relu, dropout_p, seed = inputs
    _y, mean, rstd = output[:3]
    ctx.flags = (relu, training, dropout_p, seed, weight is not None, bias is not None)
    ctx.save_for_backward(x, weight, bias, mean, rstd)


def _bn_bwd(ctx, douts):
    dy = douts[0]
    x, weight, bias, mean, rstd = ctx.saved_tensors
    relu, training, dropout_p, seed, has_w, has_b = ctx.flags
    dx, dw, db = torch.ops.gnnx.bn_relu_bwd(dy, x, weight, bias, mean, rstd, relu, training, dropout_p, seed)
    return dx, (dw if has_w else None), (db if has_b else None), None, None, None, None, None, None, None, None


bn_relu.register_autograd(_bn_bwd, setup_context=_bn_setup)


# ------------------------------------------------------------------ graph pooling
@custom_op("gnnx::segment_pool", mutates_args=())
def segment_pool(x: Tensor, graph_ptr: Tensor, mode: int) -> Tensor:
    return ops.segment_pool_fwd(x.contiguous(), graph_ptr, int(mode))


@segment_pool.register_fake
def _(x, graph_ptr, mode):
    return x.new_empty((graph_ptr.shape[0] - 1, x.shape[1]))


@custom_op("gnnx::segment_pool_bwd", mutates_args=())
def segment_pool_bwd(dout: Tensor, x: Tensor, out: Tensor, graph_ptr: Tensor, mode: int) -> Tensor:
    return ops.segment_pool_bwd(dout.contiguous(), x, out, graph_ptr, int(mode))


@segment_pool_bwd.register_fake
def _(dout, x, out, graph_ptr, mode):
    return x.new_empty(x.shape)


def _pool_setup(ctx, inputs, output):
    x, graph_ptr, mode = inputs
    ctx.mode = mode
    ctx.save_for_backward(x, output, graph_ptr)


def _pool_bwd(ctx, dout):
    x, out, graph_ptr = ctx.saved_tensors
    return torch.ops.gnnx.segment_pool_bwd(dout, x, out, graph_ptr, ctx.mode), None, None


segment_pool.register_autograd(_pool_bwd, setup_context=_pool_setup)


# ------------------------------------------------------------------ conv layers (one op per layer)
def _index_from(parts: List[Tensor], num_nodes: int, max_degree: int):
    from .graph import GraphIndex
    g = GraphIndex(int(num_nodes), int(parts[1].shape[0]), tuple(parts))
    if max_degree >= 0:
        g.max_degree = int(max_degree)
    return g


def _run_pna(x, efeat, codes, index, params, towers, pre_layers, post_layers, avg_deg_log, max_degree):
    from . import pna_fn
    g = _index_from(index, x.shape[0], max_degree)
    H = x.shape[1]
    meta = pna_fn.PNAMeta(g, codes, towers, H // towers, H // towers, pre_layers, post_layers, avg_deg_log)
    if max_degree < 0 and pna_fn.grouped_ok(meta):
        # no host-side degree bound (a traced graph must not read one back): the general scaler-folded path
        return _PNAGeneral.run(x, efeat, meta, params)
    return pna_fn.PNAConvFn.apply(x, efeat, meta, *params)


class _PNAGeneral:
    """PNAConvFn with the degree-grouped path switched off for one call (module switch, restored afterwards)."""

    @staticmethod
    def run(x, efeat, meta, params):
        from . import pna_fn
        was = pna_fn.GROUPED_POST
        pna_fn.GROUPED_POST = False
        try:
            return pna_fn.PNAConvFn.apply(x, efeat, meta, *params)
        finally:
            pna_fn.GROUPED_POST = was


@custom_op("gnnx::pna_conv", mutates_args=())
def pna_conv(x: Tensor, efeat: Tensor, codes: Optional[Tensor], index: List[Tensor], params: List[Tensor], towers: int,
             pre_layers: int, post_layers: int, avg_deg_log: float, max_degree: int) -> Tensor:
    """``torch_geometric.nn.PNAConv.forward`` (models.py:445-457) on a prebuilt index.  ``efeat``: edge features
    [E, edge_dim] (``codes`` None) or the bond table [60, edge_dim] with one uint8 code per dst-sorted slot;
    ``max_degree`` >= 0: a host-side bound on the in-degree (enables the degree-grouped first post Linear), -1: none."""
    with torch.no_grad():
        return _run_pna(x, efeat, codes, list(index), list(params), towers, pre_layers, post_layers, avg_deg_log, max_degree)


@pna_conv.register_fake
def _(x, efeat, codes, index, params, towers, pre_layers, post_layers, avg_deg_log, max_degree):
    return x.new_empty((x.shape[0], params[-1].shape[0]))


def _run_gat(x, efeat, codes, index, params, heads, out_channels, negative_slope, add_self_loops, dropout_p, seed):
    from . import gat_fn
    g = _index_from(index, x.shape[0], -1)
    meta = gat_fn.GATMeta(g, codes, heads, out_channels, negative_slope, add_self_loops, dropout_p, seed)
    return gat_fn.GATv2ConvFn.apply(x, efeat, meta, *params)


@custom_op("gnnx::gatv2_conv", mutates_args=())
def gatv2_conv(x: Tensor, efeat: Tensor, codes: Optional[Tensor], index: List[Tensor], params: List[Tensor], heads: int,
               out_channels: int, negative_slope: float, add_self_loops: bool, dropout_p: float, seed: int) -> Tensor:
    """``GATv2Conv.forward`` (models.py:487-495); params = lin_l.{weight,bias}, lin_r.{weight,bias}, lin_edge.weight, att, bias."""
    with torch.no_grad():
        return _run_gat(x, efeat, codes, list(index), list(params), heads, out_channels, negative_slope, add_self_loops,
                        dropout_p, seed)


@gatv2_conv.register_fake
def _(x, efeat, codes, index, params, heads, out_channels, negative_slope, add_self_loops, dropout_p, seed):
    return x.new_empty((x.shape[0], heads * out_channels))


def _run_tconv(x, efeat, codes, index, params, heads, out_channels, dropout_p, seed):
    from . import tconv_fn
    g = _index_from(index, x.shape[0], -1)
    meta = tconv_fn.TConvMeta(g, codes, heads, out_channels, dropout_p, seed)
    return tconv_fn.TransformerConvFn.apply(x, efeat, meta, *params)


@custom_op("gnnx::transformer_conv", mutates_args=())
def transformer_conv(x: Tensor, efeat: Tensor, codes: Optional[Tensor], index: List[Tensor], params: List[Tensor],
                     heads: int, out_channels: int, dropout_p: float, seed: int) -> Tensor:
    """``TransformerConv.forward`` (models.py:497-510); params = lin_query / lin_key / lin_value {weight,bias},
    lin_edge.weight, lin_skip.{weight,bias}."""
    with torch.no_grad():
        return _run_tconv(x, efeat, codes, list(index), list(params), heads, out_channels, dropout_p, seed)


@transformer_conv.register_fake
def _(x, efeat, codes, index, params, heads, out_channels, dropout_p, seed):
    return x.new_empty((x.shape[0], heads * out_channels))


def _conv_autograd(op, runner, n_static):
    """Backward by recomputation: re-run the fused autograd node with grad enabled on detached inputs and pull the
    gradients of (x, efeat, params) out of it.  Dropout draws are functions of (seed, index), so the replay is exact."""

    def setup(ctx, inputs, output):
        x, efeat, codes, index, params = inputs[:5]
        ctx.static = inputs[5:]
        ctx.n_index, ctx.n_params, ctx.has_codes = len(index), len(params), codes is not None
        ctx.save_for_backward(x, efeat, *( [codes] if codes is not None else [] ), *index, *params)

    def backward(ctx, dout):
        saved = list(ctx.saved_tensors)
        x, efeat = saved[0], saved[1]
        pos = 2
        codes = saved[pos] if ctx.has_codes else None
        pos += int(ctx.has_codes)
        index = saved[pos:pos + ctx.n_index]
        params = saved[pos + ctx.n_index:]
        with torch.enable_grad():
            xg = x.detach().requires_grad_(True)
            eg = efeat.detach().requires_grad_(True)
            pg = [p.detach().requires_grad_(True) for p in params]
            out = runner(xg, eg, codes, list(index), pg, *ctx.static)
            grads = torch.autograd.grad(out, [xg, eg, *pg], dout.contiguous(), allow_unused=True)
        pgr = [g if g is not None else torch.zeros_like(p) for g, p in zip(grads[2:], params)]
        return (grads[0], grads[1], None, [None] * ctx.n_index, pgr) + (None,) * n_static

    op.register_autograd(backward, setup_context=setup)


_conv_autograd(pna_conv, _run_pna, 5)
_conv_autograd(gatv2_conv, _run_gat, 6)
_conv_autograd(transformer_conv, _run_tconv, 4)
