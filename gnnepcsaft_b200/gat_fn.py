"""GATv2Conv forward / backward as one autograd node over the gnnx kernels."""

from __future__ import annotations

import torch
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from . import ops
from .functional import _c


# Debug tap: when set to a list, every forward appends dict(alpha [E + N, heads] -- dst-sorted slots then loop slots,
# a negative sign marks an entry removed by attention dropout --, g) so tests can replay the draw on the oracle.
CAPTURE_ALPHA = None


class GATMeta:
    __slots__ = ("g", "codes", "heads", "c", "slope", "self_loops", "dropout_p", "seed")

    def __init__(self, g, codes, heads, c, slope, self_loops, dropout_p=0.0, seed=0):
        self.g, self.codes, self.heads, self.c, self.slope, self.self_loops = g, codes, heads, c, slope, self_loops
        self.dropout_p, self.seed = dropout_p, seed


class GATv2ConvFn(Function):
    """params: lin_l.weight, lin_l.bias, lin_r.weight, lin_r.bias, lin_edge.weight, att, bias."""

    @staticmethod
    def forward(ctx, x, efeat, meta: GATMeta, wl, bl, wr, br, wedge, att, bias):
        g, heads, c = meta.g, meta.heads, meta.c
        H = heads * c
        N = x.shape[0]
        x, efeat = _c(x), _c(efeat)
        xlr = torch.empty((N, 2 * H), dtype=torch.float32, device=x.device)
        ops.gemm(x, wl, trans_b=True, bias=bl, out=xlr[:, :H])
        ops.gemm(x, wr, trans_b=True, bias=br, out=xlr[:, H:])
        e = ops.gemm(efeat, wedge, trans_b=True)                  # [E or 60, H], lin_edge has no bias
        eidx = meta.codes if meta.codes is not None else g.csr_eid
        att_flat = att.reshape(-1)
        out, alpha = ops.gatv2_fwd(xlr, e, eidx, att_flat, bias, g, heads, c, meta.slope, meta.self_loops,
                                   meta.dropout_p, meta.seed)
        if CAPTURE_ALPHA is not None:
            CAPTURE_ALPHA.append(dict(alpha=alpha, g=g))
        ctx.meta = meta
        ctx.has = (bl is not None, br is not None, bias is not None)
        ctx.save_for_backward(x, efeat, xlr, e, alpha, wl, wr, wedge, att, out, bias)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        meta: GATMeta = ctx.meta
        g, heads, c = meta.g, meta.heads, meta.c
        H = heads * c
        x, efeat, xlr, e, alpha, wl, wr, wedge, att, out, bias = ctx.saved_tensors
        dout = _c(dout)
        eidx = meta.codes if meta.codes is not None else g.csr_eid
        table = meta.codes is not None
        # the edge-term gradient of every slot (loop fan-out included) comes out of the same pass as ds
        dxlr, datt, ds, de = ops.gatv2_bwd(dout, out, bias, xlr, e, eidx, att.reshape(-1), alpha, g, heads, c, meta.slope,
                                           meta.self_loops, meta.dropout_p, want_de="slot" if table else "edge")
        dl, dr = dxlr[:, :H], dxlr[:, H:]
        dx = dbl = dbr = None
        if ctx.needs_input_grad[0]:                                  # the bias gradients ride the dX products
            dx, dbl = ops.gemm(dl, wl, want_colsum_a=True)
            _, dbr = ops.gemm(dr, wr, out=dx, accumulate=True, want_colsum_a=True)
        else:
            dbl, dbr = ops.colsum(dl), ops.colsum(dr)
        dbl = dbl if ctx.has[0] else None
        dbr = dbr if ctx.has[1] else None
        dwl, dwr = ops.gemm_dw_batched([dl, dr], [x, x])             # two same-shape weight gradients, one launch
        dbias = ops.colsum(dout) if ctx.has[2] else None
        if table:
            de = ops.code_segment_sum(de, meta.codes, efeat.shape[0])
        dwedge = ops.gemm(de, efeat, trans_a=True)
        defeat = ops.gemm(de, wedge) if ctx.needs_input_grad[1] else None
        return dx, defeat, None, dwl, dbl, dwr, dbr, dwedge, datt.view_as(att), dbias
