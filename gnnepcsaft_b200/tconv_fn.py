"""TransformerConv forward / backward as one autograd node over the gnnx kernels.

The three node projections run as ONE product ``x [W_q; W_k; W_v]^T`` (N = 3H), the skip projection writes the
output buffer that the attention kernel then adds the aggregate to, and the bias-free edge projection is applied to
the 60-row bond table when the batch carries one (reference: gnnepcsaft/train/models.py:497-510).
"""

from __future__ import annotations

import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from . import ops
from .functional import _c


# Debug tap: when set to a list, every forward appends dict(alpha [E, heads] in dst-sorted slot order -- a negative
# sign marks an entry removed by attention dropout --, g) so that tests can replay the same dropout draw on the oracle.
CAPTURE_ALPHA = None


class TConvMeta:
    __slots__ = ("g", "codes", "heads", "c", "dropout_p", "seed")

    def __init__(self, g, codes, heads, c, dropout_p=0.0, seed=0):
        self.g, self.codes, self.heads, self.c = g, codes, heads, c
        self.dropout_p, self.seed = dropout_p, seed


class TransformerConvFn(Function):
    """params: lin_query.{weight,bias}, lin_key.{weight,bias}, lin_value.{weight,bias}, lin_edge.weight,
    lin_skip.{weight,bias}."""

    @staticmethod
    def forward(ctx, x, efeat, meta: TConvMeta, wq, bq, wk, bk, wv, bv, wedge, wskip, bskip):
        g, heads, c = meta.g, meta.heads, meta.c
        x, efeat = _c(x), _c(efeat)
        wqkv = torch.cat([wq, wk, wv], 0)
        bqkv = torch.cat([bq, bk, bv], 0)
        qkv = ops.gemm(x, wqkv, trans_b=True, bias=bqkv)              # [N, 3H]
        e = ops.gemm(efeat, wedge, trans_b=True)                     # [E or 60, H]
        out = ops.gemm(x, wskip, trans_b=True, bias=bskip)           # lin_skip(x); the kernel adds the aggregate
        eidx = meta.codes if meta.codes is not None else g.csr_eid
        agg, alpha = ops.tconv_fwd(qkv, e, eidx, g, heads, c, out, meta.dropout_p, meta.seed)
        if CAPTURE_ALPHA is not None:
            CAPTURE_ALPHA.append(dict(alpha=alpha, g=g))
        ctx.meta = meta
        ctx.save_for_backward(x, efeat, qkv, e, alpha, agg, wqkv, wedge, wskip)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        meta: TConvMeta = ctx.meta
        g, heads, c = meta.g, meta.heads, meta.c
        H = heads * c
        x, efeat, qkv, e, alpha, agg, wqkv, wedge, wskip = ctx.saved_tensors
        dout = _c(dout)
        table = meta.codes is not None
        eidx = meta.codes if table else g.csr_eid
        dqkv, de = ops.tconv_bwd(dout, qkv, e, eidx, alpha, agg, g, heads, c, original_order=not table,
                                 dropout_p=meta.dropout_p)
        if ctx.needs_input_grad[0]:                                  # the bias gradients ride the dX products
            dx, dbqkv = ops.gemm(dqkv, wqkv, want_colsum_a=True)
            _, dbskip = ops.gemm(dout, wskip, out=dx, accumulate=True, want_colsum_a=True)
        else:
            dx, dbqkv, dbskip = None, ops.colsum(dqkv), ops.colsum(dout)
        # four same-shape weight gradients (q, k, v, skip), one batched launch
        dwq, dwk, dwv, dwskip = ops.gemm_dw_batched([dqkv[:, :H], dqkv[:, H:2 * H], dqkv[:, 2 * H:], dout], [x, x, x, x])
        if table:
            de = ops.code_segment_sum(de, meta.codes, efeat.shape[0])
        dwedge = ops.gemm(de, efeat, trans_a=True)
        defeat = ops.gemm(de, wedge) if ctx.needs_input_grad[1] else None
        return (dx, defeat, None, dwq, dbqkv[:H], dwk, dbqkv[H:2 * H], dwv, dbqkv[2 * H:],
                dwedge, dwskip, dbskip)
