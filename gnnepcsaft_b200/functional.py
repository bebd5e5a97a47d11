"""Autograd glue: each Function strings gnnx kernels together for one reference module's
forward and backward.  Fat ops on purpose -- a conv layer is ONE autograd node, so the Python
dispatch cost per layer is a handful of ctypes calls instead of ~80 ATen ops.
"""

from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from . import ops
from .graph import GraphIndex


def _c(t: Tensor) -> Tensor:
    return t if t.is_contiguous() else t.contiguous()


# ------------------------------------------------------------------ Linear (+ReLU)
class _LinearFn(Function):
    """y = act(x W^T + b)  (torch.nn.Linear semantics; the MLP head, models.py:186-194)."""

    @staticmethod
    def forward(ctx, x, weight, bias, relu):
        y = ops.gemm(x, weight, trans_b=True, bias=bias, relu=relu)
        ctx.relu = relu
        ctx.save_for_backward(x, weight, y if relu else None)
        ctx.has_bias = bias is not None
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        dy = _c(dy)
        if ctx.relu:
            dy = ops.relu_bwd(dy, y)
        dx = ops.gemm(dy, weight) if ctx.needs_input_grad[0] else None
        dw = ops.gemm(dy, x, trans_a=True)
        db = ops.colsum(dy) if ctx.has_bias else None
        return dx, dw, db, None


def linear(x: Tensor, weight: Tensor, bias: Optional[Tensor], relu: bool = False) -> Tensor:
    return _LinearFn.apply(_c(x), weight, bias, relu)


# ------------------------------------------------------------------ BatchNorm (+ReLU)
class _BatchNormFn(Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, training, momentum, eps, relu, buckets=None,
                dropout_p=0.0, seed=0):
        y, mean, rstd = ops.bn_relu_fwd(x, weight, bias, running_mean, running_var, training, momentum, eps, relu, buckets,
                                        dropout_p, seed)
        ctx.relu, ctx.training, ctx.buckets, ctx.drop = relu, training, buckets, (dropout_p, seed)
        ctx.save_for_backward(x, weight, bias, mean, rstd)     # the backward re-derives the ReLU mask from x: y is not kept
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, weight, bias, mean, rstd = ctx.saved_tensors
        dx, dw, db = ops.bn_relu_bwd(_c(dy), x, None, weight, mean, rstd, ctx.relu, ctx.buckets, bias=bias,
                                     training=ctx.training, dropout_p=ctx.drop[0], seed=ctx.drop[1])
        return dx, dw, db, None, None, None, None, None, None, None, None, None


def batch_norm(x: Tensor, bn: torch.nn.BatchNorm1d, relu: bool = False, buckets=None, dropout_p: float = 0.0,
               seed: int = 0) -> Tensor:
    """BatchNorm1d (+ fused ReLU, + the next layer's node-feature Dropout) with the module's parameters / running
    statistics.  ``buckets``: ``x`` is a PNAConv output in degree-bucket order (``PNAConv(..., bucket_out=True)``); the
    result is in node order.  Gradients flow in eval mode too (frozen statistics: fine-tuning, saliency)."""
    training = bn.training or not bn.track_running_stats
    if training and bn.track_running_stats and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
    if bn.momentum is None:
        # torch: cumulative moving average, factor 1 / num_batches_tracked (one host read; no reference config uses it)
        momentum = 1.0 / max(1, int(bn.num_batches_tracked.item())) if (training and bn.track_running_stats) else 0.0
    else:
        momentum = bn.momentum
    return _BatchNormFn.apply(_c(x), bn.weight, bn.bias, bn.running_mean, bn.running_var, training, momentum,
                              bn.eps, relu, buckets, dropout_p, seed)


# ------------------------------------------------------------------ graph pooling
class _PoolFn(Function):
    @staticmethod
    def forward(ctx, x, graph_ptr, mode):
        out = ops.segment_pool_fwd(x, graph_ptr, mode)
        ctx.mode = mode
        ctx.save_for_backward(x, out, graph_ptr)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        x, out, graph_ptr = ctx.saved_tensors
        return ops.segment_pool_bwd(_c(dout), x, out, graph_ptr, ctx.mode), None, None


def segment_pool(x: Tensor, graph_ptr: Tensor, mode: str) -> Tensor:
    return _PoolFn.apply(_c(x), graph_ptr, ops.POOL_MODES[mode])


# ------------------------------------------------------------------ OGB embedding sums
class _EmbedFn(Function):
    @staticmethod
    def forward(ctx, table, idx_u8, offsets, dropout_p=0.0, seed=0):
        ctx.save_for_backward(idx_u8, offsets)
        ctx.rows, ctx.drop = table.shape[0], (dropout_p, seed)
        return ops.embed_sum_fwd(idx_u8, offsets, table, dropout_p, seed)

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        idx_u8, offsets = ctx.saved_tensors
        return ops.embed_sum_bwd(idx_u8, offsets, ctx.rows, _c(dout), ctx.drop[0], ctx.drop[1]), None, None, None, None


def embed_sum(table: Tensor, idx_u8: Tensor, offsets: Tensor, dropout_p: float = 0.0, seed: int = 0) -> Tensor:
    return _EmbedFn.apply(table, idx_u8, offsets, dropout_p, seed)


class _MatFn(Function):
    """out = P @ B for a constant 0/1 selector P (bond-type table = P [60,13] @ cat(bond tables))."""

    @staticmethod
    def forward(ctx, p, b):
        ctx.save_for_backward(p)
        return ops.gemm(p, b, precision=ops.PREC_FP32)

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        (p,) = ctx.saved_tensors
        return None, ops.gemm(p, _c(dout), trans_a=True, precision=ops.PREC_FP32)


def select_rows_matmul(p: Tensor, b: Tensor) -> Tensor:
    return _MatFn.apply(p, b)


class EdgeTable:
    """Edge features in factored form: ``table[codes[slot]]`` with a <=60-row table (every OGB
    bond-feature combination) and one uint8 code per dst-sorted slot.  Lets the conv layers run
    their edge Linear on 60 rows instead of E rows."""

    __slots__ = ("table", "codes")

    def __init__(self, table: Tensor, codes: Tensor):
        self.table, self.codes = table, codes
