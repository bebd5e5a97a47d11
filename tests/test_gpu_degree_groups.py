"""Degree-grouped first post_nn Linear of PNAConv (csrc/degree_groups.cu + the grouped GEMM entry points).

Integer work (bucketing, split table) is checked bit-exactly against numpy; the grouped products against an fp64
product of the same operands -- the fp32 tier to rel 1e-5 of the row scale, the tf32 tier to the same
3 * 2^-11 * (|A| |B|) bound as tests/test_gpu_gemm_tc.py.
"""
import numpy as np
import pytest
import torch

from gnnepcsaft_b200 import graph, ops
from gnnepcsaft_b200._lib import GnnxError

pytestmark = pytest.mark.gpu
DEV = "cuda"
U_TF32 = 2.0 ** -11


def _rowptr(deg):
    return torch.from_numpy(np.concatenate([[0], np.cumsum(deg)]).astype(np.int32)).to(DEV)


def _buckets(deg, n_groups=None):
    n_groups = int(deg.max()) + 1 if n_groups is None else n_groups
    return graph.DegreeBuckets(_rowptr(deg), n_groups)


@pytest.mark.parametrize("N,maxdeg", [(1, 0), (5, 4), (800, 4), (1000, 0), (4097, 6), (101887, 4), (70000, 15)])
def test_degree_buckets_bit_exact(N, maxdeg):
    rng = np.random.default_rng(N)
    deg = rng.integers(0, maxdeg + 1, N)
    deg[rng.integers(0, N)] = maxdeg
    if N > 100:
        deg[deg == 1] = 2                                   # an empty bucket in the middle
    bk = _buckets(deg)
    D, A = maxdeg + 1, ops.GROUP_ALIGN
    counts = np.bincount(deg, minlength=D)
    bp = np.concatenate([[0], np.cumsum((counts + A - 1) // A * A)])
    assert bk.bucket_ptr.cpu().numpy().tolist() == bp.tolist()
    assert int(bk.err_flag.item()) == 0
    order = np.argsort(deg, kind="stable")                   # stable: node order inside a bucket
    pos = np.empty(N, np.int64)
    start = 0
    for d in range(D):
        pos[order[start:start + counts[d]]] = bp[d] + np.arange(counts[d])
        start += counts[d]
    assert np.array_equal(bk.node_pos.cpu().numpy(), pos)
    row_node = np.full(bk.capacity, -1, np.int64)
    row_node[pos] = np.arange(N)
    assert np.array_equal(bk.row_node.cpu().numpy(), row_node)
    assert bk.capacity >= bp[-1] and bk.capacity % A == 0


def test_degree_buckets_overflow_raises_flag():
    deg = np.array([0, 1, 2, 9, 3])
    bk = _buckets(deg, n_groups=4)
    assert int(bk.err_flag.item()) == 1
    with pytest.raises(ValueError):
        _buckets(deg, n_groups=ops.MAX_GROUPS + 1)


@pytest.mark.parametrize("splits", [5, 29, 74, 148])
def test_split_table_covers_every_bucket_once(splits):
    rng = np.random.default_rng(splits)
    deg = rng.choice(5, 50000, p=[0.0, 0.35, 0.37, 0.25, 0.03])
    bk = _buckets(deg, n_groups=5)
    bp = bk.bucket_ptr.cpu().numpy()
    t = bk.split_table(splits).cpu().numpy()
    covered = np.zeros(bp[-1], np.int32)
    for r0, r1, g in t:
        if g < 0:
            assert r0 == r1
            continue
        assert bp[g] <= r0 <= r1 <= bp[g + 1] and r0 % 32 == 0 and r1 % 32 == 0
        covered[r0:r1] += 1
    assert np.all(covered == 1)
    used = t[t[:, 2] >= 0]
    assert len(used) >= min(splits, 4) and (np.diff(used[:, 0]) >= 0).all()
    # balanced: no split is more than ~2x the ideal share
    assert (used[:, 1] - used[:, 0]).max() <= 2 * max(32, bp[-1] / splits) + 32


def _grouped_ref(a64, w64, bp, D, trans_b):
    out = None
    for g in range(D):
        r0, r1 = int(bp[g]), int(bp[g + 1]) if g < D - 1 else a64.shape[0]
        wg = w64[g]
        o = a64[r0:r1] @ (wg.t() if trans_b else wg)
        out = o if out is None else torch.cat([out, o])
    return out


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
@pytest.mark.parametrize("N,K,No,trans_b", [(3000, 320, 64, True), (3000, 64, 320, False), (20000, 1280, 256, True),
                                            (20000, 256, 1280, False), (700, 160, 96, True)])
def test_gemm_grouped(prec, N, K, No, trans_b):
    torch.manual_seed(1)
    ops.set_matmul_precision(prec)
    try:
        rng = np.random.default_rng(7)
        deg = rng.choice(5, N, p=[0.01, 0.35, 0.36, 0.25, 0.03])
        bk = _buckets(deg, n_groups=5)
        D, M = 5, bk.capacity
        bp = bk.bucket_ptr.cpu().numpy()
        a = torch.randn(M, K)
        w = torch.randn(D, No, K) if trans_b else torch.randn(D, K, No)
        bias = torch.randn(No)
        out = ops.gemm_grouped(a.to(DEV), w.reshape(-1, w.shape[2]).to(DEV), bk, trans_b=trans_b, bias=bias.to(DEV), relu=True)
        ref = _grouped_ref(a.double(), w.double(), bp, D, trans_b) + bias.double()
        absab = _grouped_ref(a.double().abs(), w.double().abs(), bp, D, trans_b)
        bound = (3 * U_TF32 if prec == "tf32" else 1e-5) * absab + 1e-6
        err = (out.double().cpu() - ref.clamp(min=0)).abs()
        assert float((err / bound).max()) <= 1.0, f"grouped product error/bound {float((err / bound).max()):.3g}"
        out2, cs = ops.gemm_grouped(a.to(DEV), w.reshape(-1, w.shape[2]).to(DEV), bk, trans_b=trans_b, want_colsum_a=True)
        assert float(((out2.double().cpu() - _grouped_ref(a.double(), w.double(), bp, D, trans_b)).abs() / bound).max()) <= 1.0
        assert torch.allclose(cs.cpu().double(), a.double().sum(0), rtol=1e-4, atol=1e-4 * float(a.abs().sum(0).max()))
        with pytest.raises((ValueError, GnnxError)):
            ops.gemm_grouped(a[:-1].to(DEV), w.reshape(-1, w.shape[2]).to(DEV), bk, trans_b=trans_b)
    finally:
        ops.set_matmul_precision("fp32")


@pytest.mark.parametrize("prec", ["fp32", "tf32"])
@pytest.mark.parametrize("N,Mo,No", [(3000, 64, 320), (30000, 256, 1280), (900, 96, 160)])
def test_gemm_grouped_dw(prec, N, Mo, No):
    torch.manual_seed(2)
    ops.set_matmul_precision(prec)
    try:
        rng = np.random.default_rng(11)
        deg = rng.choice(5, N, p=[0.0, 0.35, 0.37, 0.25, 0.03])
        bk = _buckets(deg, n_groups=5)
        bp = bk.bucket_ptr.cpu().numpy()
        R = bk.capacity
        a, b = torch.randn(R, Mo), torch.randn(R, No)
        valid = (bk.row_node >= 0).cpu()
        a[~valid] = 0                                           # padding rows carry zeros in the product path
        coef = torch.rand(5, 3) + 0.5
        out = ops.gemm_grouped_dw(a.to(DEV), b.to(DEV), bk, coef.to(DEV))
        a64, b64 = a.double(), b.double()
        ref = torch.zeros(3, Mo, No, dtype=torch.float64)
        absab = torch.zeros(Mo, No, dtype=torch.float64)
        for g in range(5):
            r0, r1 = int(bp[g]), int(bp[g + 1])
            p = a64[r0:r1].t() @ b64[r0:r1]
            absab += a64[r0:r1].abs().t() @ b64[r0:r1].abs()
            for j in range(3):
                ref[j] += float(coef[g, j]) * p
        bound = (3 * U_TF32 if prec == "tf32" else 2e-5) * 1.5 * absab + 1e-5
        err = (out.double().cpu() - ref).abs()
        assert float((err / bound).max()) <= 1.0, f"grouped dW error/bound {float((err / bound).max()):.3g}"
        again = ops.gemm_grouped_dw(a.to(DEV), b.to(DEV), bk, coef.to(DEV))
        assert torch.equal(out, again), "grouped dW is not deterministic"
    finally:
        ops.set_matmul_precision("fp32")


def test_gather_rows_and_bits():
    torch.manual_seed(3)
    src = torch.randn(500, 320, device=DEV)
    idx = torch.randint(-1, 500, (777,), device=DEV, dtype=torch.int32)
    out, bits = ops.gather_rows(src[:, 32:], idx, cols=96, want_relu_bits=True)
    ref = torch.where((idx >= 0)[:, None], src[idx.clamp(min=0).long(), 32:128], torch.zeros((), device=DEV))
    assert torch.equal(out, ref)
    b = bits.cpu().numpy().astype(np.uint32)
    pos = (ref > 0).cpu().numpy()
    for w in range(3):
        want = (pos[:, 32 * w:32 * w + 32].astype(np.uint64) << np.arange(32, dtype=np.uint64)).sum(1).astype(np.uint32)
        assert np.array_equal(b[:, w], want)


def test_degree_weights():
    torch.manual_seed(4)
    Ho, H, D, avg = 64, 96, 6, 0.9
    wz, wx = torch.randn(3 * Ho, 4 * H), torch.randn(Ho, H)
    weff, coef = ops.pna_degree_weights(wz.to(DEV), wx.to(DEV), avg, D)
    d = torch.arange(D, dtype=torch.float32)
    sa, st = torch.log(d + 1) / avg, avg / torch.log(d.clamp(min=1) + 1)
    assert torch.allclose(coef.cpu(), torch.stack([torch.ones(D), sa, st], 1), rtol=1e-6, atol=1e-7)
    ref = torch.cat([wz[None, :Ho] + sa[:, None, None] * wz[None, Ho:2 * Ho] + st[:, None, None] * wz[None, 2 * Ho:],
                     wx[None].expand(D, -1, -1)], 2)
    assert torch.allclose(weff.cpu().view(D, Ho, 5 * H), ref, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("N,C,relu", [(5000, 256, True), (777, 64, False)])
def test_batchnorm_on_bucket_ordered_rows(N, C, relu):
    """BatchNorm fed in degree-bucket order == BatchNorm on the same rows in node order (statistics are sums over the
    same rows in another order: fp32 tolerance 1e-5), padding rows excluded, their gradient exactly zero."""
    from gnnepcsaft_b200 import functional as Fn
    torch.manual_seed(5)
    rng = np.random.default_rng(5)
    deg = rng.choice(5, N, p=[0.02, 0.35, 0.35, 0.25, 0.03])
    bk = _buckets(deg, n_groups=5)
    x = torch.randn(N, C, device=DEV) * 2 + 0.5
    xb = torch.full((bk.capacity, C), 7.0, device=DEV)               # padding rows hold finite junk
    xb[bk.node_pos.long()] = x
    outs = []
    for inp, rows in ((x, None), (xb, bk)):
        bn = torch.nn.BatchNorm1d(C).to(DEV).train()
        torch.manual_seed(6)                                         # the same affine parameters for both runs
        with torch.no_grad():
            bn.weight.uniform_(0.5, 1.5); bn.bias.uniform_(-0.5, 0.5)
        inp = inp.clone().requires_grad_(True)
        y = Fn.batch_norm(inp, bn, relu, rows)
        w = torch.linspace(-1, 1, N * C, device=DEV).view(N, C)
        (y * w).sum().backward()
        outs.append((y.detach(), inp.grad, bn.weight.grad, bn.bias.grad, bn.running_mean.clone(), bn.running_var.clone()))
    (y0, dx0, dw0, db0, rm0, rv0), (y1, dx1, dw1, db1, rm1, rv1) = outs
    assert y1.shape == (N, C)
    for a, b, what in ((y1, y0, "y"), (dx1[bk.node_pos.long()], dx0, "dx"), (dw1, dw0, "dweight"), (db1, db0, "dbias"),
                       (rm1, rm0, "running_mean"), (rv1, rv0, "running_var")):
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-5 * float(b.abs().max())), what
    pad = (bk.row_node < 0)
    assert bool((dx1[pad] == 0).all())


def test_degree_hint_skips_the_readback_and_flags_a_batch_that_breaks_it():
    """Reference-signature forward of a PNA model: the dataset histogram's length bounds the in-degree, so no per-batch
    maximum is read back; a batch with a node beyond the histogram still runs (wrongly bucketed), is flagged on the
    device, and the NEXT forward raises.  With the hint switched off the same batch is served by one read-back."""
    import copy
    import gnnepcsaft_b200 as G
    from tests.util import PNA_CFG1, deg_hist, make_batch
    ops.set_matmul_precision("fp32")
    graphs, b = make_batch(16, seed=5)
    hist = deg_hist(graphs)
    args = [b[k].to(DEV) for k in ("x", "edge_index", "edge_attr", "batch")]
    m = G.GNNePCSAFT(copy.deepcopy(dict(PNA_CFG1, deg=hist))).to(DEV).eval()
    assert m.degree_hint == len(hist) - 1
    with torch.no_grad():
        y_hint = m(*args, num_graphs=16)
        idx = graph.graph_index_for(args[1], args[0].shape[0])
        assert idx.max_degree is None and idx.degree_hint == m.degree_hint and idx._status_host is not None
        m.check_pending()                                            # nothing flagged
        m.degree_hint = None
        graph._INDEX_CACHE.clear()
        y_read = m(*args, num_graphs=16)
        assert torch.equal(y_hint, y_read)
        # a model whose histogram stops one short of this batch's largest in-degree
        short = G.GNNePCSAFT(copy.deepcopy(dict(PNA_CFG1, deg=hist[:-1]))).to(DEV).eval()
        graph._INDEX_CACHE.clear()
        short(*args, num_graphs=16)                                  # runs, flagged on the device
        with pytest.raises(IndexError, match="degree histogram"):
            short(*args, num_graphs=16)
        short.degree_hint = None
        graph._INDEX_CACHE.clear()
        assert bool(torch.isfinite(short(*args, num_graphs=16)).all())
