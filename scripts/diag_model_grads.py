"""Debug aid: whole-model training step on the CUDA path vs the fp32 / fp64 oracle, one line per parameter gradient.

    python scripts/diag_model_grads.py [pna|gat]
"""
import copy
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import gnnepcsaft_b200 as G                                     # noqa: E402
from gnnepcsaft_b200 import ops, pna_fn                          # noqa: E402
from oracle import pyg_restated as O                             # noqa: E402
from tests.util import GAT_CFG_SMALL, PNA_CFG1, branches_from, deg_hist, make_batch  # noqa: E402


WORST = {}


def _note(kind, info, err, scale):
    rel = err / max(scale, 1e-30)
    if kind not in WORST or rel > WORST[kind][0]:
        WORST[kind] = (rel, info)
    if rel > 3e-5:
        print(f"  !! {kind} {info}: err {err:.3g} scale {scale:.3g} rel {rel:.3g}")


def install_op_checks():
    """Wrap gemm / colsum / bn_relu_bwd so every call is compared with an fp64 torch evaluation of the same op."""
    g0, c0, b0 = ops.gemm, ops.colsum, ops.bn_relu_bwd

    def gemm(a, b, *, trans_a=False, trans_b=False, bias=None, relu=False, out=None, accumulate=False, precision=None,
             relu_mask=None):
        prev = out.double().clone() if accumulate else None
        res = g0(a, b, trans_a=trans_a, trans_b=trans_b, bias=bias, relu=relu, out=out, accumulate=accumulate,
                 precision=precision, relu_mask=relu_mask)
        A, B = a.double(), b.double()
        ref = (A.t() if trans_a else A) @ (B.t() if trans_b else B)
        absref = (A.t() if trans_a else A).abs() @ (B.t() if trans_b else B).abs()
        if bias is not None:
            ref = ref + bias.double()
        if prev is not None:
            ref = ref + prev
        if relu_mask is not None:
            ref = ref * (relu_mask > 0)
        if relu:
            ref = ref.clamp_min(0)
        _note("gemm", f"M{res.shape[0]} N{res.shape[1]} K{A.shape[0] if trans_a else A.shape[1]} ta{int(trans_a)} tb{int(trans_b)} "
              f"mask{int(relu_mask is not None)} acc{int(accumulate)}", float((res.double() - ref).abs().max()), float(absref.max()))
        return res

    def colsum(x, out=None):
        res = c0(x, out)
        ref = x.double().sum(0)
        _note("colsum", f"{tuple(x.shape)} ld{x.stride(0)}", float((res.double() - ref).abs().max()), float(x.double().abs().sum(0).max()))
        return res

    def bn_relu_bwd(dy, x, y, weight, mean, rstd, relu, buckets=None, bias=None):
        dx, dw, db = b0(dy, x, y, weight, mean, rstd, relu, buckets, bias=bias)
        if y is None:
            y = (x - mean) * rstd * weight + (bias if bias is not None else 0.0)
        g = dy.double() * (y > 0) if relu else dy.double()
        xh = (x.double() - mean.double()) * rstd.double()
        rdb, rdw = g.sum(0), (g * xh).sum(0)
        n = x.shape[0]
        rdx = weight.double() * rstd.double() * (g - rdb / n - xh * rdw / n)
        sc = float(g.abs().sum(0).max())
        _note("bn_bwd.db", f"{tuple(x.shape)}", float((db.double() - rdb).abs().max()), sc)
        _note("bn_bwd.dw", f"{tuple(x.shape)}", float((dw.double() - rdw).abs().max()), float((g * xh).abs().sum(0).max()))
        _note("bn_bwd.dx", f"{tuple(x.shape)}", float((dx.double() - rdx).abs().max()), float(rdx.abs().max()))
        return dx, dw, db

    ops.gemm, ops.colsum, ops.bn_relu_bwd = gemm, colsum, bn_relu_bwd


def main():
    install_op_checks()
    which = sys.argv[1] if len(sys.argv) > 1 else "pna"
    cfg, target = (PNA_CFG1, "para") if which == "pna" else (GAT_CFG_SMALL, "assoc")
    torch.set_float32_matmul_precision("highest")
    ops.set_matmul_precision("fp32")
    torch.manual_seed(0)
    graphs, b = make_batch(32, seed=1234)
    cfg = dict(cfg, deg=deg_hist(graphs))
    torch.manual_seed(1)
    mo = O.GNNePCSAFT(cfg)
    mp = G.GNNePCSAFT(copy.deepcopy(cfg)).to("cuda")
    mp.load_state_dict(mo.state_dict(), strict=True)
    mo64 = copy.deepcopy(mo).double()
    mo.train(); mp.train(); mo64.train()
    taps = {}

    def tap(model, tag):
        def mk(name):
            def hook(_m, _i, out):
                taps[(tag, name, "out")] = out.detach().double().cpu()
                out.register_hook(lambda gr: taps.__setitem__((tag, name, "grad"), gr.detach().double().cpu()))
            return hook
        for i, c in enumerate(model.convs):
            c.register_forward_hook(mk(f"conv{i}"))
        for i, c in enumerate(model.batch_norms):
            c.register_forward_hook(mk(f"bn{i}"))
    tap(mo, "o32"); tap(mo64, "o64"); tap(mp, "gpu")
    pna_fn.CAPTURE_AGG = []
    pp = mp(b["x"].cuda(), b["edge_index"].cuda(), b["edge_attr"].cuda(), b["batch"].cuda())
    aggs = list(pna_fn.CAPTURE_AGG)
    pna_fn.CAPTURE_AGG = None
    relu_out = [taps[("gpu", f"bn{i}", "out")] for i in range(len(mp.batch_norms))]
    with branches_from(aggs, cfg["towers"], relu_out):
        po = mo(b["x"], b["edge_index"], b["edge_attr"], b["batch"])
    with branches_from(aggs, cfg["towers"], relu_out):
        p64 = mo64(b["x"], b["edge_index"], b["edge_attr"], b["batch"])
    print("fwd err gpu", float((pp.cpu().double() - p64).abs().max()), "ref", float((po.double() - p64).abs().max()))
    lo, _ = O.training_loss(po, b[target])
    l64, _ = O.training_loss(p64, b[target].double())
    lo.backward(); l64.backward()
    _, dpred = ops.relhuber_fwd_bwd(pp.detach().contiguous(), b[target].cuda())
    pp.backward(dpred)
    for (tag, name, kind), v in sorted(taps.items()):
        if tag != "gpu":
            continue
        r32, r64 = taps[("o32", name, kind)], taps[("o64", name, kind)]
        if name.startswith("bn") and kind == "out":
            r32, r64 = r32.clamp_min(0), r64.clamp_min(0)
        if name.startswith("bn") and kind == "grad":   # post-ReLU (gpu) vs pre-ReLU (oracle): apply the mask
            v = v * (taps[("gpu", name, "out")] > 0)
        print(f"tap {name:6s} {kind:4s} scale {float(r64.abs().max()):9.3g} err_gpu {float((v - r64).abs().max()):9.3g} "
              f"err_ref {float((r32 - r64).abs().max()):9.3g}")
    for i in range(3):
        gg = taps[("gpu", f"bn{i}", "grad")] * (taps[("gpu", f"bn{i}", "out")] > 0)
        d = (gg - taps[("o64", f"bn{i}", "grad")]).abs()
        for idx in torch.nonzero(d > 1e-9)[:8]:
            r, c = int(idx[0]), int(idx[1])
            print(f"bn{i} grad mismatch at ({r},{c}): gpu dy {float(taps[('gpu', f'bn{i}', 'grad')][r, c]):.6g} "
                  f"y_gpu {float(taps[('gpu', f'bn{i}', 'out')][r, c]):.9g} y32 {float(taps[('o32', f'bn{i}', 'out')][r, c]):.9g} "
                  f"y64 {float(taps[('o64', f'bn{i}', 'out')][r, c]):.9g} g64 {float(taps[('o64', f'bn{i}', 'grad')][r, c]):.6g} "
                  f"x_gpu {float(taps[('gpu', f'conv{i}', 'out')][r, c]):.9g} x64 {float(taps[('o64', f'conv{i}', 'out')][r, c]):.9g}")
        print(f"bn{i}: {int((d > 1e-9).sum())} mismatching entries")
    po_, po64 = dict(mo.named_parameters()), dict(mo64.named_parameters())
    for n, p in mp.named_parameters():
        g, r32, r64 = p.grad.cpu().double(), po_[n].grad.double(), po64[n].grad
        eg, er, sc = float((g - r64).abs().max()), float((r32 - r64).abs().max()), float(r64.abs().max())
        flag = "  <<<<" if eg > max(4 * er, 2e-6 * sc) else ""
        if flag or os.environ.get("DIAG_ALL"):
            print(f"{n:50s} shape {str(tuple(g.shape)):14s} scale {sc:9.3g} err_gpu {eg:9.3g} err_ref {er:9.3g}{flag}")


if __name__ == "__main__":
    main()
    for k, (rel, info) in WORST.items():
        print(f"worst {k}: rel {rel:.3g} at {info}")

