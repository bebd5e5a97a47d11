"""PNAConv forward / backward as one autograd node over the gnnx kernels.

Restructurings relative to the PyG formulation, all algebraically exact:
  * first pre_nn Linear split node-side:  W[x_i | x_j | e'] = (x W_i^T)[dst] + (x W_j^T)[src] + e' W_e^T,
    so the [E, 3F] concat is never built and two of the three products run on N rows, not E;
  * the four aggregators are produced by ONE pass over the dst-sorted messages;
  * the degree scalers are per-row scalars, so (s * agg) W^T = s * (agg W^T): the [N, 12F] / [N, 13F]
    concats are never built -- ONE product agg [N,4H] x Wz^T [4H, 3Ho] plus a fused combine replace the K=13F product;
  * towers are folded into block-diagonal weights (built per call from the per-tower parameters, a few KB..MB):
    every Linear of the layer is then ONE full-width GEMM whatever `towers` is.  These products are HBM-bound at
    the hot path's shapes, so the zero blocks cost tensor-pipe flops, not time;
  * the ReLU backward between Linear layers rides the dX product's epilogue (`relu_mask`), fed by a 1-bit-per-element
    mask that the forward product (or the gather kernel) packed from its own output;
  * the last post_nn Linear has no activation before `lin`, so the two are applied as ONE Linear with
    Wf = W_lin W_last, bf = W_lin b_last + b_lin (a 256^3 weight-space product per call); the chain rule back to
    the two original parameters is exact (dW_lin = dWf W_last^T + dbf b_last^T, dW_last = W_lin^T dWf, ...).
"""

from __future__ import annotations

from typing import List

import torch
from torch import Tensor
from torch.autograd import Function
from torch.autograd.function import once_differentiable

from . import ops
from .functional import _c


# Degree-grouped first post Linear (include/gnnx.h, csrc/degree_groups.cu): used whenever the batch has at most
# ops.MAX_GROUPS in-degree classes; False forces the general scaler-folded path (tests compare the two).
GROUPED_POST = True

# Debug taps: when set to a list, every PNAConv forward appends dict(agg [N, T*4*F], m [E, T*F] dst-sorted, g,
# pre_relu [..], post_relu [..]) (tests read which Std entries were zero-masked, which messages were the min / max of
# their row, and the sign pattern of every ReLU inside the layer).
CAPTURE_AGG = None
CAPTURE_BWD = None


class PNAMeta:
    """Non-tensor arguments of one PNAConv call."""

    __slots__ = ("g", "codes", "towers", "f_in", "f_out", "pre_layers", "post_layers", "avg_deg_log", "prep", "bucket_out")

    def __init__(self, g, codes, towers, f_in, f_out, pre_layers, post_layers, avg_deg_log, prep=None):
        self.g, self.codes, self.towers, self.f_in, self.f_out = g, codes, towers, f_in, f_out
        self.pre_layers, self.post_layers, self.avg_deg_log = pre_layers, post_layers, avg_deg_log
        self.prep = prep              # parameter-only tensors computed ahead of the call (PNAPrep), or None
        self.bucket_out = False       # leave the output (and take dout) in degree-bucket order


def _split_params(meta: PNAMeta, params):
    it = iter(params)
    we, be = next(it), next(it)
    pre = [[(next(it), next(it)) for _ in range(meta.pre_layers)] for _ in range(meta.towers)]
    post = [[(next(it), next(it)) for _ in range(meta.post_layers)] for _ in range(meta.towers)]
    wl, bl = next(it), next(it)
    return we, be, pre, post, wl, bl


def _bd(blocks: List[Tensor]) -> Tensor:
    return blocks[0].contiguous() if len(blocks) == 1 else torch.block_diag(*blocks)


def _cat(ts: List[Tensor], dim: int = 0) -> Tensor:
    return ts[0].contiguous() if len(ts) == 1 else torch.cat(ts, dim)


class _Packed:
    """Block-diagonal / stacked views of the per-tower parameters (see module docstring)."""

    def __init__(self, meta: PNAMeta, pre, post):
        T, F = meta.towers, meta.f_in
        w0 = [pre[t][0][0] for t in range(T)]
        self.wpq = torch.cat([_bd([w[:, 0:F] for w in w0]), _bd([w[:, F:2 * F] for w in w0])], 0)      # [2H, H]
        self.wr = _cat([w[:, 2 * F:3 * F] for w in w0])                                                # [H, F]
        self.br = _cat([pre[t][0][1] for t in range(T)])
        self.pre = [(_bd([pre[t][l][0] for t in range(T)]), _cat([pre[t][l][1] for t in range(T)]))
                    for l in range(1, meta.pre_layers)]
        wp = [post[t][0][0] for t in range(T)]
        self.wx = _bd([w[:, 0:F] for w in wp])                                                         # [Ho, H]
        self.wz = torch.cat([_bd([w[:, F + s * 4 * F:F + (s + 1) * 4 * F] for w in wp]) for s in range(3)], 0)  # [3Ho, 4H]
        self.b0 = _cat([post[t][0][1] for t in range(T)])
        self.post = [(_bd([post[t][l][0] for t in range(T)]), _cat([post[t][l][1] for t in range(T)]))
                     for l in range(1, meta.post_layers)]


def _chain_fwd(inp: Tensor, layers, relu_last: bool = False, want_masks: bool = True):
    """Linear + ReLU layers (no ReLU after the last unless ``relu_last``); returns the input and every layer output,
    and for every output the packed ReLU mask its backward needs (``None`` where no ReLU was applied, or when no
    backward will run)."""
    outs, bits = [inp], [None]
    for i, (w, b) in enumerate(layers):
        relu = relu_last or i < len(layers) - 1
        if relu and want_masks:
            o, m = ops.gemm(outs[-1], w, trans_b=True, bias=b, relu=True, want_relu_bits=True)
        else:
            o, m = ops.gemm(outs[-1], w, trans_b=True, bias=b, relu=relu), None
        outs.append(o); bits.append(m)
    return outs, bits


def _wprod(a: Tensor, b: Tensor, **kw) -> Tensor:
    """Weight-space product (a few hundred rows): always on the fp32 kernel."""
    return ops.gemm(a, b, precision=ops.PREC_FP32, **kw)


def _chain_bwd(d: Tensor, acts: List[Tensor], masks, layers, extra=None):
    """Backward of ``_chain_fwd``.  ``acts[i]`` is the input of layer i and ``masks[i]`` the packed mask of the ReLU
    it went through (``None``: no ReLU to undo here); that ReLU's backward is fused into the dX product.
    Returns (d at acts[0], dW, db).  The weight gradients of the chain -- and of ``extra`` = (dY, X) of a layer above
    it with the same shapes -- are ONE batched product (``ops.gemm_dw_batched``) once every dY is known; with
    ``extra`` the result is (d, dW, db, dW_extra)."""
    n = len(layers)
    dys, dbs = [None] * n, [None] * n
    for i in range(n - 1, -1, -1):
        w, _ = layers[i]
        dys[i] = d
        d, dbs[i] = ops.gemm(d, w, relu_mask=masks[i], want_colsum_a=True)      # db = colsum(d) rides the dX product
    pairs = [(dys[i], acts[i]) for i in range(n)] + ([extra] if extra is not None else [])
    same = len({(a.shape, b.shape) for a, b in pairs}) == 1
    if len(pairs) >= 2 and len(pairs) <= 4 and same:
        dws = ops.gemm_dw_batched([a for a, _ in pairs], [b for _, b in pairs])
    else:
        dws = [ops.gemm(a, b, trans_a=True) for a, b in pairs]
    if extra is not None:
        return d, dws[:n], dbs, dws[n]
    return d, dws, dbs


def _diag_blocks(m: Tensor, T: int) -> List[Tensor]:
    r, c = m.shape[0] // T, m.shape[1] // T
    return [m[t * r:(t + 1) * r, t * c:(t + 1) * c] for t in range(T)]


def grouped_ok(meta: PNAMeta) -> bool:
    """Widths the degree-grouped tensor-core products take (32-column operand chunks)."""
    return GROUPED_POST and (meta.towers * meta.f_out) % 32 == 0 and (meta.towers * meta.f_in) % 32 == 0


class PNAPrep:
    """Everything a PNAConv call derives from its PARAMETERS alone (and the 60-row bond table): packed weights, the
    edge-encoder tables, the fused ``W_lin W_last`` weight and the per-degree weights.  Computed inline by the call, or
    ahead of it on the side stream (``event`` then marks its completion) -- GNNePCSAFT prepares all layers at once."""

    __slots__ = ("pk", "e1", "r", "wf", "bf", "weff", "coef", "n_groups", "event")


def prepare(meta: PNAMeta, params, table: Tensor = None, n_groups: int = 0) -> PNAPrep:
    """``table``: the bond-feature table [codes, edge_dim] when the edge features are factored (else the edge terms
    are per-edge products and stay with the call); ``n_groups`` > 0: also the per-degree weights of the grouped path."""
    we, be, pre, post, wl, bl = _split_params(meta, params)
    p = PNAPrep()
    p.pk = pk = _Packed(meta, pre, post)
    p.e1 = p.r = p.weff = p.coef = p.event = None
    p.n_groups = n_groups
    if table is not None:
        # edge term: edge_encoder, then the e' block of the first pre_nn Linear (+ its bias); 60 rows in table form
        p.e1 = ops.gemm(table, we, trans_b=True, bias=be)
        p.r = ops.gemm(p.e1, pk.wr, trans_b=True, bias=pk.br)
    if pk.post:                                       # fold the activation-free last post Linear into `lin`
        w_last, b_last = pk.post[-1]
        p.wf = _wprod(wl, w_last)
        p.bf = _wprod(wl, b_last.unsqueeze(1)).squeeze(1) + bl
    else:
        p.wf, p.bf = wl, bl
    if n_groups:
        p.weff, p.coef = ops.pna_degree_weights(pk.wz, pk.wx, meta.avg_deg_log, n_groups)
    return p


class PNAConvFn(Function):
    @staticmethod
    def forward(ctx, x, efeat, meta: PNAMeta, *params):
        g, T, F, Fo = meta.g, meta.towers, meta.f_in, meta.f_out
        Ho = T * Fo
        we, be, pre, post, wl, bl = _split_params(meta, params)
        x = _c(x)
        efeat = _c(efeat)
        bk = g.degree_buckets() if grouped_ok(meta) else None
        prep = meta.prep
        if prep is None or prep.n_groups != (bk.n_groups if bk is not None else 0):
            prep = prepare(meta, params, efeat if meta.codes is not None else None, bk.n_groups if bk is not None else 0)
        elif prep.event is not None:
            torch.cuda.current_stream().wait_event(prep.event)
        pk = prep.pk
        e1, r = prep.e1, prep.r
        if e1 is None:                                # dense edge features: per-edge products
            e1 = ops.gemm(efeat, we, trans_b=True, bias=be)
            r = ops.gemm(e1, pk.wr, trans_b=True, bias=pk.br)
        pq = ops.gemm(x, pk.wpq, trans_b=True)
        ridx = meta.codes if meta.codes is not None else g.csr_eid
        train = any(ctx.needs_input_grad)             # False under no_grad: no ReLU masks to keep
        if meta.pre_layers > 1 and train:
            h, hbits = ops.pna_gather_fwd(pq, r, ridx, g.rowptr, g.csr_src, relu=True, want_relu_bits=True)
        else:
            h, hbits = ops.pna_gather_fwd(pq, r, ridx, g.rowptr, g.csr_src, relu=meta.pre_layers > 1), None
        hs, hmasks = _chain_fwd(h, pk.pre, want_masks=train)
        hmasks[0] = hbits
        m = hs[-1]
        relu0 = meta.post_layers > 1
        obits, weff, coef = None, prep.weff, prep.coef
        if bk is not None:
            # rows bucketed by in-degree: [agg | x] W_eff(d)^T, one grouped product with a third of the flops
            fused = ops.pna_post_fused_ok(bk, T, F, Ho)
            if fused:
                # aggregate -> product in ONE tcgen05 kernel: [agg | x] is produced in shared memory, and written out
                # only when a backward (its weight gradient) or a test tap will read it
                want_bits = relu0 and train and meta.bucket_out
                o, obits, agg = ops.pna_post_fused_fwd(m, x, g, bk, T, weff, pk.b0, relu0, want_relu_bits=want_bits,
                                                       want_agg=train or CAPTURE_AGG is not None)
            else:
                agg = ops.pna_aggregate_rows_fwd(m, x, g, bk, T)                   # [cap, 5H], bucket order
            if CAPTURE_AGG is not None:
                CAPTURE_AGG.append(dict(agg=ops.gather_rows(agg, bk.node_pos, cols=4 * T * F), m=m, g=g))
            if fused:
                if not meta.bucket_out:
                    if relu0 and train:
                        o, obits = ops.gather_rows(o, bk.node_pos, want_relu_bits=True)    # back to node order
                    else:
                        o = ops.gather_rows(o, bk.node_pos)
            elif meta.bucket_out:                      # the rest of the layer is row-wise: stay in bucket order
                if relu0 and train:
                    o, obits = ops.gemm_grouped(agg, weff, bk, trans_b=True, bias=pk.b0, relu=True, want_relu_bits=True)
                else:
                    o = ops.gemm_grouped(agg, weff, bk, trans_b=True, bias=pk.b0, relu=relu0)
            else:
                o_b = ops.gemm_grouped(agg, weff, bk, trans_b=True, bias=pk.b0, relu=relu0)
                if relu0 and train:
                    o, obits = ops.gather_rows(o_b, bk.node_pos, want_relu_bits=True)   # back to node order
                else:
                    o = ops.gather_rows(o_b, bk.node_pos)
        elif meta.bucket_out:
            raise ValueError("PNAConv: bucket_out needs the degree-grouped path (bucket_rows(g) is None for this layer)")
        else:
            agg = ops.pna_aggregate_fwd(m, g.rowptr, T)
            if CAPTURE_AGG is not None:
                CAPTURE_AGG.append(dict(agg=agg, m=m, g=g))
            z = torch.empty((x.shape[0], 3 * Ho), dtype=torch.float32, device=x.device)
            ops.gemm(agg, pk.wz, trans_b=True, out=z)
            ops.gemm(x, pk.wx, trans_b=True, out=z[:, :Ho], accumulate=True)
            o = ops.pna_post_combine_fwd(z, pk.b0, g.rowptr, meta.avg_deg_log, relu=relu0)
        wf, bf = prep.wf, prep.bf
        mids = pk.post[:-1] if pk.post else []
        if mids and wf.shape[0] == wf.shape[1] == o.shape[1] and all(w.shape == wf.shape for w, _ in mids):
            # post layers 1 .. and the fused (last post Linear o lin) as ONE launch: a row tile goes through all of them
            # and its intermediates come back from L2 (ops.gemm_chain; layer-by-layer where the kernel does not apply)
            couts, cbits = ops.gemm_chain(o, [(w, b, True, train) for w, b in mids] + [(wf, bf, False, False)])
            os_, omasks, out = [o] + couts[:-1], [None] + cbits[:-1], couts[-1]
        else:
            if pk.post:
                os_, omasks = _chain_fwd(o, pk.post[:-1], relu_last=True, want_masks=train)
            else:
                os_, omasks = [o], [None]
            out = ops.gemm(os_[-1], wf, trans_b=True, bias=bf)
        if CAPTURE_AGG is not None:
            # ReLU'd activations inside the pre / post chains (tests teacher-force their sign pattern onto the oracle):
            # pre in dst-sorted slot order, post in node order
            cap = CAPTURE_AGG[-1]
            cap["pre_relu"] = list(hs[:-1]) if meta.pre_layers > 1 else []
            post_relu = list(os_) if meta.post_layers > 1 else []
            if meta.bucket_out:
                post_relu = [ops.gather_rows(t, bk.node_pos) for t in post_relu]
            cap["post_relu"] = post_relu
        ctx.meta = meta
        ctx.pk, ctx.wf = pk, (wf if pk.post else None)     # packed / fused weights of this call, reused by the backward
        ctx.n_hs, ctx.n_os = len(hs), len(os_)
        ctx.bk, ctx.weff, ctx.coef = bk, weff, coef
        omasks[0] = obits                              # general path: the combine output's own ReLU is undone by combine_bwd
        ctx.mask_slots = [m is not None for m in hmasks + omasks]
        ctx.save_for_backward(x, efeat, e1, agg, *hs, *os_, *[m for m in hmasks + omasks if m is not None], *params)
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
        pos = 4 + ctx.n_hs + ctx.n_os
        masks = []
        for present in ctx.mask_slots:
            masks.append(saved[pos] if present else None)
            pos += int(present)
        hmasks, omasks = masks[:ctx.n_hs], masks[ctx.n_hs:]
        params = saved[pos:]
        we, be, pre, post, wl, bl = _split_params(meta, params)
        pk, wf = ctx.pk, ctx.wf
        grads: dict = {}
        dout = _c(dout)

        # lin (fused with the last post Linear when there is one); its weight gradient joins the chain's batched product
        dwf = None if pk.post else ops.gemm(dout, os_[-1], trans_a=True)
        if pk.post:
            w_last, b_last = pk.post[-1]
            # os_[-1] went through a ReLU; for os_[0] (the combine output) combine_bwd undoes it
            d, dbf = ops.gemm(dout, wf, relu_mask=omasks[-1], want_colsum_a=True)
            # chain rule back to W_lin / W_last / their biases: weight-space products, queued with the side work below
        else:
            d, dbf = ops.gemm(dout, wl, want_colsum_a=True)
            grads[id(wl)], grads[id(bl)] = dwf, dbf
        # post layers 1..post-2 -> gradient at o (the combine output; its own ReLU is undone by combine_bwd)
        if pk.post:
            d, dws, dbs, dwf = _chain_bwd(d, os_, omasks, pk.post[:-1], extra=(dout, os_[-1]))
        else:
            d, dws, dbs = _chain_bwd(d, os_, omasks, pk.post[:-1])
        for l, (dw, db) in enumerate(zip(dws, dbs), start=1):
            for t, blk in enumerate(_diag_blocks(dw, T)):
                grads[id(post[t][l][0])] = blk
                grads[id(post[t][l][1])] = db[t * Fo:(t + 1) * Fo]
        bk = ctx.bk
        dxp = None
        if bk is not None:
            # degree-grouped first post Linear: d is already masked by the ReLU (packed bits rode the dX product above)
            # [cap, Ho] in bucket order, padding rows 0 (bucket_out: dout arrived that way and the chain above is row-wise)
            dyb = d if meta.bucket_out else ops.gather_rows(d, bk.row_node)
            dagg, db0 = ops.gemm_grouped(dyb, ctx.weff, bk, trans_b=False, want_colsum_a=True)   # d[agg | x]  [cap, 5H]
            gw = ops.gemm_grouped_dw(dyb, agg, bk, ctx.coef)                        # [3, Ho, 5H]
            dwz = gw[:, :, :4 * H].reshape(3 * Ho, 4 * H)
            dwx = gw[0, :, 4 * H:]
            dz = dy0 = None
        else:
            # first post Linear with folded scalers
            dz = ops.pna_post_combine_bwd(d, os_[0], g.rowptr, meta.avg_deg_log, relu=meta.post_layers > 1)
            dy0 = dz[:, :Ho]
            dwz = ops.gemm(dz, agg, trans_a=True)                   # [3Ho, 4H]
            dwx = ops.gemm(dy0, x, trans_a=True)                    # [Ho, H]
            db0 = ops.colsum(dy0)
            dagg = ops.gemm(dz, pk.wz)                              # [N, 4H]
        for t in range(T):
            cols = [dwx[t * Fo:(t + 1) * Fo, t * F:(t + 1) * F]]
            cols += [dwz[s * Ho + t * Fo:s * Ho + (t + 1) * Fo, t * 4 * F:(t + 1) * 4 * F] for s in range(3)]
            grads[id(post[t][0][0])] = torch.cat(cols, 1)
            grads[id(post[t][0][1])] = db0[t * Fo:(t + 1) * Fo]
        # aggregators
        if bk is not None:
            dm, dxp = ops.pna_aggregate_rows_bwd(dagg, hs[-1], agg, g, bk, T)
        else:
            dm = ops.pna_aggregate_bwd(dagg, hs[-1], agg, g.rowptr, T)
        if CAPTURE_BWD is not None:
            CAPTURE_BWD.append(dict(dagg=dagg, dm=dm, m=hs[-1], agg=agg, dz=dz, d_post=d))
        # pre layers 1.. -> gradient at the gather output (ReLU of the first layer fused into the last dX product)
        dh, dws, dbs = _chain_bwd(dm, hs, hmasks, pk.pre)
        for l, (dw, db) in enumerate(zip(dws, dbs), start=1):
            for t, blk in enumerate(_diag_blocks(dw, T)):
                grads[id(pre[t][l][0])] = blk
                grads[id(pre[t][l][1])] = db[t * F:(t + 1) * F]
        table = meta.codes is not None
        dpq, dr = ops.pna_gather_bwd(dh, g, want_dr_edge=not table)
        if table:
            dr = ops.code_segment_sum(dh, meta.codes, efeat.shape[0])
        dx = None
        if ctx.needs_input_grad[0]:
            if dxp is not None:                                 # the x block of d[agg | x], already in node order
                dx = ops.gemm(dpq, pk.wpq, out=dxp, accumulate=True)
            else:
                dx = ops.gemm(dpq, pk.wpq)
                ops.gemm(dy0, pk.wx, out=dx, accumulate=True)
        dwpq = ops.gemm(dpq, x, trans_a=True)                   # [2H, H]
        # ---- weight-space tail (parameters and 60-row tables only)
        if pk.post:                                             # chain rule of the fused lin back to W_lin / W_last / biases
            w_last, b_last = pk.post[-1]
            grads[id(wl)] = _wprod(dwf, w_last, trans_b=True) + torch.outer(dbf, b_last)
            grads[id(bl)] = dbf
            dw_last = _wprod(wl, dwf, trans_a=True)
            db_last = _wprod(wl, dbf.unsqueeze(1), trans_a=True).squeeze(1)
            l_last = meta.post_layers - 1
            for t, blk in enumerate(_diag_blocks(dw_last, T)):
                grads[id(post[t][l_last][0])] = blk
                grads[id(post[t][l_last][1])] = db_last[t * Fo:(t + 1) * Fo]
        dwr = ops.gemm(dr, e1, trans_a=True)                    # [H, F]
        dbr = ops.colsum(dr)
        de1 = ops.gemm(dr, pk.wr)
        grads[id(we)] = ops.gemm(de1, efeat, trans_a=True)
        grads[id(be)] = ops.colsum(de1)
        defeat = ops.gemm(de1, we) if ctx.needs_input_grad[1] else None
        for t in range(T):
            sl = slice(t * F, (t + 1) * F)
            grads[id(pre[t][0][0])] = torch.cat([dwpq[sl, sl], dwpq[H + t * F:H + (t + 1) * F, sl], dwr[sl]], 1)
            grads[id(pre[t][0][1])] = dbr[sl]
        return (dx, defeat, None, *[grads.get(id(p)) for p in params])
