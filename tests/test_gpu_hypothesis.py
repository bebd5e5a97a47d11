"""Property-based tests (hypothesis) of the integer collation work and the segmented reductions, as SURVEY sections 4 /
7.2 ask: random multigraphs with zero-edge graphs, isolated nodes, self loops, duplicate edges and hub rows.

CSR / degree / bucket work must be BIT-EXACT against the CPU oracle; the fused multi-aggregator is compared with the
oracle's scatter-based mean / min / max (exact: pure selections and a short ordered sum) and std (rel 1e-5 away from the
Std threshold, where the oracle's branch is a coin flip of the last bit).
"""
import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

from gnnepcsaft_b200 import graph, ops
from oracle import pyg_restated as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
SETTINGS = dict(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])


@st.composite
def multigraphs(draw, max_nodes=120, max_edges=600):
    n = draw(st.integers(1, max_nodes))
    e = draw(st.integers(0, max_edges))
    seed = draw(st.integers(0, 2 ** 31 - 1))
    hub = draw(st.booleans())
    rng = np.random.default_rng(seed)
    ei = rng.integers(0, n, size=(2, e))
    if hub and e > 8:
        ei[1, : e // 2] = rng.integers(0, n)          # one destination with half of the edges
    return n, torch.from_numpy(ei.astype(np.int64))


@settings(**SETTINGS)
@given(multigraphs())
def test_csr_is_the_stable_argsort_of_the_destinations(case):
    n, ei = case
    g = graph.GraphIndex.build(ei.to(DEV), n, check=True)
    ref = O.csr_reference(ei, n)
    assert torch.equal(g.rowptr.cpu().long(), ref["rowptr"])
    assert torch.equal(g.csr_eid.cpu().long(), ref["eid"])
    assert torch.equal(g.csr_src.cpu().long(), ref["src"])
    assert torch.equal(g.colptr.cpu().long(), ref["colptr"])
    assert torch.equal(g.csc_dst.cpu().long(), ref["dst_by_src"])
    assert torch.equal(g.in_degree.cpu().long(), ref["in_degree"])
    # the permutation is a permutation, and it is stable: equal destinations keep their input order
    eid = g.csr_eid.cpu().long()
    assert sorted(eid.tolist()) == list(range(ei.shape[1]))
    dst_sorted = ei[1][eid]
    assert bool((dst_sorted[1:] >= dst_sorted[:-1]).all())
    same = dst_sorted[1:] == dst_sorted[:-1]
    assert bool((eid[1:][same] > eid[:-1][same]).all())


@settings(**SETTINGS)
@given(multigraphs(max_nodes=300, max_edges=900))
def test_degree_histogram_and_buckets(case):
    n, ei = case
    g = graph.GraphIndex.build(ei.to(DEV), n)
    deg = torch.bincount(ei[1], minlength=n)
    hist, mx = graph.calc_deg_from_index(g, nbins=1024)
    m = int(mx.item())
    assert m == int(deg.max()) and hist[: m + 1].tolist() == torch.bincount(deg, minlength=m + 1).tolist()
    if m >= ops.MAX_GROUPS:
        assert g.degree_buckets() is None
        return
    bk = g.degree_buckets()
    pos, row_node, bp = bk.node_pos.cpu().long(), bk.row_node.cpu().long(), bk.bucket_ptr.cpu().long()
    assert int(bk.err_flag.item()) == 0 and bk.capacity % ops.GROUP_ALIGN == 0
    for d in range(m + 1):                               # bucket d holds exactly the nodes of in-degree d, in node order
        rows = row_node[bp[d]:bp[d + 1]]
        nodes = rows[rows >= 0]
        assert nodes.tolist() == torch.nonzero(deg == d).flatten().tolist()
        assert (bp[d + 1] - bp[d]) % ops.GROUP_ALIGN == 0
    assert torch.equal(row_node[pos], torch.arange(n))


@settings(**SETTINGS)
@given(multigraphs(max_nodes=80, max_edges=400), st.sampled_from([4, 32, 36, 128]))
def test_fused_multi_aggregator_against_scatter(case, H):
    n, ei = case
    E = ei.shape[1]
    g = graph.GraphIndex.build(ei.to(DEV), n)
    gen = torch.Generator().manual_seed(E * 131 + n)
    m = torch.randn(E, H, generator=gen)
    if E > 3:
        m[2] = m[0]                                       # exact ties
    dst = ei[1][g.csr_eid.cpu().long()]                   # messages live in dst-sorted slot order
    agg = ops.pna_aggregate_fwd(m.to(DEV), g.rowptr, 1).cpu().view(n, 4, H)
    mean = O.scatter(m, dst, n, "mean")
    assert torch.allclose(agg[:, 0], mean, rtol=1e-6, atol=1e-6)
    assert torch.equal(agg[:, 1], O.scatter(m, dst, n, "min"))
    assert torch.equal(agg[:, 2], O.scatter(m, dst, n, "max"))
    # std = sqrt(E[x^2] - E[x]^2) in fp32 (PyG's formula, R4): the subtraction carries ~eps * E[x^2] of absolute error
    # whatever the variance is, so the comparison with the fp64 oracle is made on the VARIANCE with that bound
    var64 = O.var_aggregation(m.double(), dst, n)
    ex2 = O.scatter(m.double() * m.double(), dst, n, "mean")
    away = (var64 - 1e-5).abs() > 4e-7 * ex2 + 1e-7         # the Std mask is a step at var = 1e-5
    got_var = agg[:, 3].double() ** 2
    want_var = torch.where(var64 > 1e-5, var64, torch.zeros_like(var64))
    err = (got_var - want_var).abs()
    assert bool((err[away] <= 1e-5 * want_var[away] + 4e-7 * ex2[away] + 1e-12).all())
    empty = torch.bincount(ei[1], minlength=n) == 0
    assert float(agg[empty].abs().sum()) == 0.0           # empty rows: 0 for all four


# ------------------------------------------------------------------ kernels with grid barriers / cross-item dependencies
FEW = dict(max_examples=15, deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])


@settings(**FEW)
@given(st.integers(8, 4000), st.integers(1, 80), st.booleans(), st.integers(0, 2 ** 31 - 1))
def test_cooperative_batchnorm_against_fp64(rows, c4, relu, seed):
    """k_bn_fwd_coop / k_bn_bwd_coop on arbitrary row counts (grids from 1 CTA to the residency cap) and channel counts
    (partial and multiple 256-column slabs): outputs, input gradient and parameter gradients against an fp64 evaluation
    of the same formulas, with the ReLU mask taken from the kernel's own output sign."""
    from gnnepcsaft_b200 import nn as gnn
    C = 4 * c4
    gen = torch.Generator().manual_seed(seed)
    x = (torch.randn(rows, C, generator=gen) * 2 + 0.5).to(DEV).requires_grad_(True)
    bn = gnn.BatchNorm(C).to(DEV).train()
    with torch.no_grad():
        bn.module.weight.uniform_(0.5, 1.5); bn.module.bias.normal_()
    y = bn(x, relu=relu)
    w = torch.randn(rows, C, generator=gen).to(DEV)
    (y * w).sum().backward()
    xd, wd, bd = x.detach().double(), bn.module.weight.detach().double(), bn.module.bias.detach().double()
    mu, var = xd.mean(0), xd.var(0, unbiased=False)
    xh = (xd - mu) / torch.sqrt(var + bn.module.eps)
    pre = xh * wd + bd
    keep = (y.detach() > 0).double() if relu else torch.ones_like(pre)
    want = pre * keep
    scale = float(want.abs().max()) + 1e-6
    assert float((y.detach().double() - want).abs().max()) <= 5e-5 * scale + 1e-5
    g = w.double() * keep
    db, dw = g.sum(0), (g * xh).sum(0)
    dx = (wd / torch.sqrt(var + bn.module.eps)) * (g - db / rows - xh * dw / rows)
    tol = 1e-4 * (float(dx.abs().max()) + 1e-6) + 1e-5
    assert float((x.grad.double() - dx).abs().max()) <= tol
    assert float((bn.module.bias.grad.double() - db).abs().max()) <= 1e-4 * (float(g.abs().sum(0).max()) + 1e-6)
    assert float((bn.module.weight.grad.double() - dw).abs().max()) <= 1e-4 * (float((g * xh).abs().sum(0).max()) + 1e-6)


@settings(**FEW)
@given(st.integers(768, 9000), st.sampled_from([32, 64, 96, 128, 256]), st.integers(2, 4), st.integers(0, 2 ** 31 - 1))
def test_linear_chain_equals_layer_by_layer(M, W, n, seed):
    ops.set_matmul_precision("tf32")
    ops.FORCE_TC = True
    try:
        gen = torch.Generator().manual_seed(seed)
        x = torch.randn(M, W, generator=gen).to(DEV)
        layers = [((torch.randn(W, W, generator=gen) / W ** 0.5).to(DEV), (torch.randn(W, generator=gen) * 0.1).to(DEV), p < n - 1, p < n - 1)
                  for p in range(n)]
        outs, bits = ops.gemm_chain(x, layers)
        a = x
        for p, (w, b, relu, want) in enumerate(layers):
            if want:
                ref, rbits = ops.gemm(a, w, trans_b=True, bias=b, relu=relu, want_relu_bits=True)
                assert torch.equal(bits[p], rbits)
            else:
                ref = ops.gemm(a, w, trans_b=True, bias=b, relu=relu)
            assert torch.equal(outs[p], ref), f"layer {p} of {n}, M={M}, W={W}"
            a = ref
    finally:
        ops.FORCE_TC = False
        ops.set_matmul_precision("fp32")


@settings(**FEW)
@given(st.integers(300, 60000), st.integers(1, 16), st.integers(1, 12), st.integers(0, 2 ** 31 - 1))
def test_weight_gradient_fold_equals_separate_reduce(R, mo32, no32, seed):
    import os
    ops.set_matmul_precision("tf32")
    ops.FORCE_TC = True
    was = os.environ.get("GNNX_RED_FOLD")
    try:
        gen = torch.Generator().manual_seed(seed)
        a = torch.randn(R, 32 * mo32, generator=gen).to(DEV)
        b = torch.randn(R, 32 * no32, generator=gen).to(DEV)
        got = []
        for fold in ("0", "1"):
            os.environ["GNNX_RED_FOLD"] = fold
            got.append(ops.gemm(a, b, trans_a=True))
        assert torch.equal(*got), f"R={R}, Mo={32 * mo32}, No={32 * no32}"
    finally:
        ops.FORCE_TC = False
        ops.set_matmul_precision("fp32")
        if was is None:
            os.environ.pop("GNNX_RED_FOLD", None)
        else:
            os.environ["GNNX_RED_FOLD"] = was
