"""k_pna_post_fused (gnnx_pna_post_fused_fwd): PNAConv's aggregate pass and its degree-grouped first post Linear as ONE
tcgen05 kernel whose A operand is produced in shared memory from the messages.

Checked against the two-kernel path it replaces (gnnx_pna_aggregate_rows_fwd + gnnx_gemm_grouped on the tensor cores):
the kept operand copy ``agg`` must be bit-identical (same operation sequence), the product must agree with an fp64
product of that operand within the tf32 bound 3 * 2^-11 * (|A| |B|) of tests/test_gpu_gemm_tc.py, the packed ReLU mask
must be the sign pattern of the output, and the result must not depend on whether the copy is kept.
"""
import numpy as np
import pytest
import torch

from gnnepcsaft_b200 import graph, ops
from gnnepcsaft_b200._lib import lib

pytestmark = pytest.mark.gpu
DEV = "cuda"
U_TF32 = 2.0 ** -11


@pytest.fixture(autouse=True)
def _tier():
    was = ops.FUSED_POST
    ops.set_matmul_precision("tf32")
    ops.FORCE_TC = True
    ops.FUSED_POST = True            # opt-in path (slower than the two kernels it replaces: see ops.py); tested all the same
    torch.manual_seed(0)
    yield
    ops.FORCE_TC = False
    ops.FUSED_POST = was
    ops.set_matmul_precision("fp32")


class _G:
    """The two fields of GraphIndex the aggregate entry points read."""

    def __init__(self, rowptr, n):
        self.rowptr, self.num_nodes = rowptr, n


def _case(N, maxdeg, H, seed, p=None):
    rng = np.random.default_rng(seed)
    deg = rng.integers(0, maxdeg + 1, N) if p is None else rng.choice(len(p), N, p=p)
    deg[rng.integers(0, N)] = maxdeg if p is None else len(p) - 1
    rowptr = torch.from_numpy(np.concatenate([[0], np.cumsum(deg)]).astype(np.int32)).to(DEV)
    E = int(deg.sum())
    n_groups = int(deg.max()) + 1
    bk = graph.DegreeBuckets(rowptr, n_groups)
    g = torch.Generator().manual_seed(seed)
    m = torch.randn(E, H, generator=g)
    # exact ties (min == max rows) and rows whose variance sits below the Std threshold, as molecules produce them
    if E > 8:
        m[1] = m[0]
    x = torch.randn(N, H, generator=g)
    return _G(rowptr, N), bk, m.to(DEV), x.to(DEV), deg


@pytest.mark.parametrize("N,maxdeg,H,T,Ho,relu", [
    (700, 4, 64, 1, 64, True), (3000, 4, 128, 2, 128, True), (5000, 6, 256, 1, 256, True), (2500, 3, 128, 4, 64, False),
    (20000, 4, 256, 1, 256, True), (1200, 9, 64, 2, 48, True), (900, 0, 64, 1, 64, True)])
def test_fused_matches_aggregate_then_grouped_product(N, maxdeg, H, T, Ho, relu):
    g, bk, m, x, deg = _case(N, maxdeg, H, seed=N + H)
    F = H // T
    assert ops.pna_post_fused_ok(bk, T, F, Ho), "the shape must be taken by the fused kernel for this test to mean anything"
    D = bk.n_groups
    weff = (torch.randn(D * Ho, 5 * H) / (5 * H) ** 0.5).to(DEV)
    bias = torch.randn(Ho).to(DEV)
    agg_ref = ops.pna_aggregate_rows_fwd(m, x, g, bk, T)
    tc0 = lib().tc_launch_count()
    out, bits, agg = ops.pna_post_fused_fwd(m, x, g, bk, T, weff, bias, relu, want_relu_bits=True, want_agg=True)
    out2, _, none = ops.pna_post_fused_fwd(m, x, g, bk, T, weff, bias, relu)
    assert lib().tc_launch_count() - tc0 == 2 and none is None
    assert torch.equal(agg, agg_ref), "the kept operand must be what gnnx_pna_aggregate_rows_fwd writes, bit for bit"
    assert torch.equal(out, out2), "the product must not depend on whether the operand copy is kept"
    # fp64 product of the same operand, group by group
    bp = bk.bucket_ptr.cpu().numpy()
    a64, w64 = agg_ref.double().cpu(), weff.double().cpu().reshape(D, Ho, 5 * H)
    ref = torch.zeros(bk.capacity, Ho, dtype=torch.float64)
    absab = torch.zeros_like(ref)
    for d in range(D):
        r0, r1 = int(bp[d]), int(bp[d + 1])
        ref[r0:r1] = a64[r0:r1] @ w64[d].t()
        absab[r0:r1] = a64[r0:r1].abs() @ w64[d].abs().t()
    ref += bias.double().cpu()
    pre = ref.clone()
    if relu:
        ref = ref.clamp(min=0)
    tol = 3 * U_TF32 * absab + 1e-6
    err = (out.double().cpu() - ref).abs()
    valid = torch.from_numpy(bk.row_node.cpu().numpy() >= 0)
    assert bool((err[valid] <= tol[valid]).all()), f"worst excess {float((err - tol)[valid].max()):.3e}"
    # padding rows: zero operand -> act(bias)
    pad = ~valid[: int(bp[-1])]
    if pad.any():
        want = bias.cpu().clamp(min=0) if relu else bias.cpu()
        assert torch.equal(out.cpu()[: int(bp[-1])][pad], want.expand(int(pad.sum()), Ho))
    # packed mask = sign pattern of the output (entries within the tolerance of zero may fall either way: use `out` itself)
    words = bits.cpu().numpy().view(np.uint32)
    o = out.cpu().numpy()
    got = ((words[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(bk.capacity, -1)[:, :Ho].astype(bool)
    assert np.array_equal(got, o > 0)
    # and against the two-kernel tensor-core path: same operand, same weights, same tf32 products -> tf32-level agreement
    two = ops.gemm_grouped(agg_ref, weff, bk, trans_b=True, bias=bias, relu=relu)
    d2 = (two.double().cpu() - out.double().cpu()).abs()
    assert bool((d2[valid] <= 2 * tol[valid]).all())
    print(f"fused vs two-kernel: max|d| {float(d2.max()):.3e}, bit-identical: {bool(torch.equal(two, out))}; "
          f"pre-activation scale {float(pre.abs().max()):.3g}")


def test_fused_is_used_by_the_layer_and_agrees_with_the_unfused_layer():
    """PNAConv forward + backward on the tf32 tier with the fused kernel on and off: same predictions and gradients to
    tf32 rounding (the backward is the same code either way and reads the same kept operand)."""
    import copy
    import gnnepcsaft_b200 as G
    from tests.util import deg_hist, make_batch
    ops.FORCE_TC = False          # whole model: the head's [G, 3] product is not a tensor-core shape; the layers are large enough
    graphs, b = make_batch(256, seed=77)
    cfg = dict(conv="PNA", hidden_dim=128, propagation_depth=2, dropout=0.0, global_pool="add", num_para=3, towers=2,
               pre_layers=2, post_layers=3, heads=2, add_self_loops=True, model="gnn", deg=deg_hist(graphs))
    model = G.GNNePCSAFT(copy.deepcopy(cfg)).to(DEV).train()
    args = [b[k].to(DEV) for k in ("x", "edge_index", "edge_attr", "batch")]
    outs = {}
    for fused in (True, False):
        ops.FUSED_POST = fused
        try:
            model.zero_grad(set_to_none=True)
            tc0 = lib().tc_launch_count()
            p = model(*args)
            n_fwd = lib().tc_launch_count() - tc0
            p.square().mean().backward()
            outs[fused] = (p.detach().clone(), [q.grad.detach().clone() for q in model.parameters()], n_fwd)
        finally:
            ops.FUSED_POST = True    # the fixture restores the default afterwards
    (p1, g1, n1), (p0, g0, n0) = outs[True], outs[False]
    assert n1 == n0, "one fused tcgen05 launch replaces one grouped tcgen05 launch per layer"
    scale = float(p0.abs().max())
    assert float((p1 - p0).abs().max()) <= 20 * U_TF32 * scale
    # gradients: the two forwards differ by the K order of one product per layer (~1e-6), which moves a few ReLU / min /
    # max decisions of this piecewise-linear model; parity of every gradient against the oracle is test_gpu_tf32_model's
    # job (it runs through the fused kernel), this one guards against a wrong operand reaching the backward
    # (bound relative to the largest gradient of the model: training-mode BatchNorm amplifies a rounding difference in a
    # near-constant column up to 1 / sqrt(eps) = 316 times, which shows on the tensors with tiny gradients)
    gmax = max(float(c.abs().max()) for c in g0)
    for a, c in zip(g1, g0):
        assert float((a - c).abs().max()) <= 0.05 * gmax
