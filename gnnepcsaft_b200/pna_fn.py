"""PNAConv forward / backward as one autograd node over the gnnx kernels.

Restructurings relative to the PyG formulation, all algebraically exact:
  * first pre_nn Linear split node-side:  W[x_i | x_j | e'] = (x W_i^T)[dst] + (x W_j^T)[src] + e' W_e^T,
    so the [E, 3F] concat is never built and two of the three products run on N rows, not E;
  * the four aggregators are produced by ONE pass over the dst-sorted messages;
  * the degree scalers are per-row scalars, so (s * agg) W^T = s * (agg W^T): the [N, 12F] / [N, 13F]
    concats are never built -- three K=4F products plus a fused combine replace the K=13F product.
"""

from __future__ import annotations

from typing import List

import torch
from torch import Tensor
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from . import ops
from .functional import _c


# Debug taps: when set to a list, every PNAConv forward appends dict(agg [N, T*4*F], m [E, T*F] dst-sorted, g)
# (tests read which Std entries were zero-masked and which messages were the min / max of their row).
CAPTURE_AGG = None
CAPTURE_BWD = None


class PNAMeta:
    """Non-tensor arguments of one PNAConv call."""

    __slots__ = ("g", "codes", "towers", "f_in", "f_out", "pre_layers", "post_layers", "avg_deg_log")

    def __init__(self, g, codes, towers, f_in, f_out, pre_layers, post_layers, avg_deg_log):
        self.g, self.codes, self.towers, self.f_in, self.f_out = g, codes, towers, f_in, f_out
        self.pre_layers, self.post_layers, self.avg_deg_log = pre_layers, post_layers, avg_deg_log


def _split_params(meta: PNAMeta, params):
    it = iter(params)
    we, be = next(it), next(it)
    pre = [[(next(it), next(it)) for _ in range(meta.pre_layers)] for _ in range(meta.towers)]
    post = [[(next(it), next(it)) for _ in range(meta.post_layers)] for _ in range(meta.towers)]
    wl, bl = next(it), next(it)
    return we, be, pre, post, wl, bl


def _tower_chain_fwd(inp: Tensor, layers, towers: int, f: int, first: int) -> List[Tensor]:
    """Per-tower Linear(+ReLU) layers ``first..`` applied to column blocks of ``inp``; returns all outputs."""
    outs = [inp]
    n_layers = len(layers[0])
    for l in range(first, n_layers):
        cur = outs[-1]
        nxt = torch.empty_like(cur)
        for t in range(towers):
            w, b = layers[t][l]
            ops.gemm(cur[:, t * f:(t + 1) * f], w, trans_b=True, bias=b, relu=l < n_layers - 1,
                     out=nxt[:, t * f:(t + 1) * f])
        outs.append(nxt)
    return outs


def _tower_chain_bwd(d_last: Tensor, acts: List[Tensor], layers, towers: int, f: int, first: int, grads: dict):
    """Backward of ``_tower_chain_fwd``; ``d_last`` is the gradient at the last (un-activated) output.
    Returns the gradient at ``acts[0]`` (before any mask for acts[0]'s own activation)."""
    n_layers = len(layers[0])
    d = d_last
    for l in range(n_layers - 1, first - 1, -1):
        y_in = acts[l - first]
        if l < n_layers - 1:                       # this layer's output went through ReLU
            d = ops.relu_bwd(d, acts[l - first + 1])
        d_in = torch.empty_like(y_in)
        for t in range(towers):
            w, b = layers[t][l]
            sl = slice(t * f, (t + 1) * f)
            ops.gemm(d[:, sl], w, out=d_in[:, sl])
            grads[id(w)] = ops.gemm(d[:, sl], y_in[:, sl], trans_a=True)
            grads[id(b)] = ops.colsum(d[:, sl])
        d = d_in
    return d


class PNAConvFn(Function):
    @staticmethod
    def forward(ctx, x, efeat, meta: PNAMeta, *params):
        g, T, F, Fo = meta.g, meta.towers, meta.f_in, meta.f_out
        H, Ho = T * F, T * Fo
        N = x.shape[0]
        we, be, pre, post, wl, bl = _split_params(meta, params)
        x = _c(x)
        efeat = _c(efeat)
        dev = x.device
        # edge term: edge_encoder, then the e' block of the first pre_nn Linear (+ its bias)
        e1 = ops.gemm(efeat, we, trans_b=True, bias=be)
        pq = torch.empty((N, 2 * H), dtype=torch.float32, device=dev)
        r = torch.empty((efeat.shape[0], H), dtype=torch.float32, device=dev)
        for t in range(T):
            w0, b0 = pre[t][0]
            xt = x[:, t * F:(t + 1) * F]
            ops.gemm(xt, w0[:, 0:F], trans_b=True, out=pq[:, t * F:(t + 1) * F])
            ops.gemm(xt, w0[:, F:2 * F], trans_b=True, out=pq[:, H + t * F:H + (t + 1) * F])
            ops.gemm(e1, w0[:, 2 * F:3 * F], trans_b=True, bias=b0, out=r[:, t * F:(t + 1) * F])
        ridx = meta.codes if meta.codes is not None else g.csr_eid
        h = ops.pna_gather_fwd(pq, r, ridx, g.rowptr, g.csr_src, relu=meta.pre_layers > 1)
        hs = _tower_chain_fwd(h, pre, T, F, 1)
        m = hs[-1]
        agg = ops.pna_aggregate_fwd(m, g.rowptr, T)
        if CAPTURE_AGG is not None:
            CAPTURE_AGG.append(dict(agg=agg, m=m, g=g))
        y0 = torch.empty((N, Ho), dtype=torch.float32, device=dev)
        z = torch.empty((N, 3 * Ho), dtype=torch.float32, device=dev)
        for t in range(T):
            wp, _ = post[t][0]
            ops.gemm(x[:, t * F:(t + 1) * F], wp[:, 0:F], trans_b=True, out=y0[:, t * Fo:(t + 1) * Fo])
            a_t = agg[:, t * 4 * F:(t + 1) * 4 * F]
            for s in range(3):
                ops.gemm(a_t, wp[:, F + s * 4 * F:F + (s + 1) * 4 * F], trans_b=True,
                         out=z[:, (3 * t + s) * Fo:(3 * t + s + 1) * Fo])
        bias0 = post[0][0][1] if T == 1 else torch.cat([post[t][0][1] for t in range(T)])
        o = ops.pna_post_combine_fwd(y0, z, bias0, g.rowptr, meta.avg_deg_log, T, relu=meta.post_layers > 1)
        os_ = _tower_chain_fwd(o, post, T, Fo, 1)
        out = ops.gemm(os_[-1], wl, trans_b=True, bias=bl)
        ctx.meta = meta
        ctx.n_hs, ctx.n_os = len(hs), len(os_)
        ctx.save_for_backward(x, efeat, e1, agg, *hs, *os_, *params)
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        meta: PNAMeta = ctx.meta
        g, T, F, Fo = meta.g, meta.towers, meta.f_in, meta.f_out
        H, Ho = T * F, T * Fo
        saved = ctx.saved_tensors
        x, efeat, e1, agg = saved[:4]
        hs = list(saved[4:4 + ctx.n_hs])
        os_ = list(saved[4 + ctx.n_hs:4 + ctx.n_hs + ctx.n_os])
        params = saved[4 + ctx.n_hs + ctx.n_os:]
        we, be, pre, post, wl, bl = _split_params(meta, params)
        grads: dict = {}
        dout = _c(dout)
        N = x.shape[0]
        dev = x.device

        # lin
        d = ops.gemm(dout, wl)
        grads[id(wl)] = ops.gemm(dout, os_[-1], trans_a=True)
        grads[id(bl)] = ops.colsum(dout)
        # post layers 1.. (reverse) -> gradient at o (post-combine output)
        d = _tower_chain_bwd(d, os_, post, T, Fo, 1, grads)
        # first post Linear with folded scalers
        dy0, dz = ops.pna_post_combine_bwd(d, os_[0], g.rowptr, meta.avg_deg_log, T, relu=meta.post_layers > 1)
        db0 = ops.colsum(dy0)
        dx = torch.empty_like(x)
        dagg = torch.empty_like(agg)
        for t in range(T):
            wp, bp = post[t][0]
            dwp = torch.empty_like(wp)
            xt = x[:, t * F:(t + 1) * F]
            dy0_t = dy0[:, t * Fo:(t + 1) * Fo]
            ops.gemm(dy0_t, wp[:, 0:F], out=dx[:, t * F:(t + 1) * F])
            ops.gemm(dy0_t, xt, trans_a=True, out=dwp[:, 0:F])
            a_t = agg[:, t * 4 * F:(t + 1) * 4 * F]
            for s in range(3):
                dz_ts = dz[:, (3 * t + s) * Fo:(3 * t + s + 1) * Fo]
                w_s = wp[:, F + s * 4 * F:F + (s + 1) * 4 * F]
                ops.gemm(dz_ts, w_s, out=dagg[:, t * 4 * F:(t + 1) * 4 * F], accumulate=s > 0)
                ops.gemm(dz_ts, a_t, trans_a=True, out=dwp[:, F + s * 4 * F:F + (s + 1) * 4 * F])
            grads[id(wp)] = dwp
            grads[id(bp)] = db0[t * Fo:(t + 1) * Fo] if T > 1 else db0
        # aggregators
        dm = ops.pna_aggregate_bwd(dagg, hs[-1], agg, g.rowptr, T)
        if CAPTURE_BWD is not None:
            CAPTURE_BWD.append(dict(dagg=dagg, dm=dm, m=hs[-1], agg=agg, dz=dz, dy0=dy0, d_post=d))
        # pre layers 1.. (reverse) -> gradient at h (after the first layer's activation)
        dh = _tower_chain_bwd(dm, hs, pre, T, F, 1, grads)
        if meta.pre_layers > 1:
            dh = ops.relu_bwd(dh, hs[0])
        table = meta.codes is not None
        dpq, dr = ops.pna_gather_bwd(dh, g, want_dr_edge=not table)
        if table:
            dr = ops.code_segment_sum(dh, meta.codes, efeat.shape[0])
        de1 = torch.empty_like(e1)
        for t in range(T):
            w0, b0 = pre[t][0]
            dw0 = torch.empty_like(w0)
            xt = x[:, t * F:(t + 1) * F]
            dp_t = dpq[:, t * F:(t + 1) * F]
            dq_t = dpq[:, H + t * F:H + (t + 1) * F]
            dr_t = dr[:, t * F:(t + 1) * F]
            ops.gemm(dp_t, w0[:, 0:F], out=dx[:, t * F:(t + 1) * F], accumulate=True)
            ops.gemm(dq_t, w0[:, F:2 * F], out=dx[:, t * F:(t + 1) * F], accumulate=True)
            ops.gemm(dp_t, xt, trans_a=True, out=dw0[:, 0:F])
            ops.gemm(dq_t, xt, trans_a=True, out=dw0[:, F:2 * F])
            ops.gemm(dr_t, e1, trans_a=True, out=dw0[:, 2 * F:3 * F])
            ops.gemm(dr_t, w0[:, 2 * F:3 * F], out=de1, accumulate=t > 0)
            grads[id(w0)] = dw0
            grads[id(b0)] = ops.colsum(dr_t)
        grads[id(we)] = ops.gemm(de1, efeat, trans_a=True)
        grads[id(be)] = ops.colsum(de1)
        defeat = ops.gemm(de1, we) if ctx.needs_input_grad[1] else None
        return (dx if ctx.needs_input_grad[0] else None, defeat, None, *[grads.get(id(p)) for p in params])
