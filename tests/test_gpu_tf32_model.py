"""Whole-model parity of the BENCHMARKED tier: tcgen05 kind::tf32 products against the fp64 oracle.

The fp32 tier (CUDA-core GEMMs) is held to rel 1e-5 / 1e-4 in tests/test_gpu_model.py.  The tensor-core tier
rounds both operands of every product to tf32 (unit roundoff u = 2^-11) and accumulates in fp32; it is what
bench.py measures, and the reference itself trains under ``torch.set_float32_matmul_precision("medium")``
(gnnepcsaft/train/train.py:54), i.e. with coarser products still.  This file pins that tier at model level:

  * forward predictions, the loss and EVERY parameter gradient of a training step are compared with the fp64 oracle;
  * the bound is stated per tensor and derived from the tier's arithmetic, not fitted to a direction:
        max|got - ref64| <= C * u * D * max(max|ref64|, floor)          (normwise, per tensor)
    with D the number of dependent products on the longest path to that quantity (forward: D_fwd; a gradient has
    crossed the forward and the backward: 2 * D_fwd) and C = 2 (two rounded operands per product).  For gradient
    tensors the floor is 5 % of the largest gradient entry of the model (biases in front of a BatchNorm have a true
    gradient of zero and are judged on that floor);
  * every discontinuity of the path is teacher-forced to the CUDA run's branch -- the Std zero-mask, the min / max
    routing, the ReLU behind each BatchNorm, and (new for this tier) every ReLU inside the pre_nn / post_nn chains
    and the MLP head -- and each forced entry is asserted to lie within the same C * u * D_fwd band of its kink
    (tests/util.py: branches_from(..., inner_relus=True, relu_band=...)).  Without that, a tf32 run flips ~1e-3 of
    the near-zero pre-activations and the comparison measures the flips, not the arithmetic;
  * the run must actually have launched tcgen05 kernels (``gnnx_tc_launch_count``) where the shapes are large enough
    to leave the small-product fp32 kernel.

Configs: BASELINE cfg1 (hidden 64, 3 layers, 32 graphs; its products are below the tensor-core kernel's size
threshold, so this case pins the tier's dispatch and fused epilogues on small shapes) and hidden 256 / 6 layers /
256 graphs (N ~ 6.4 k rows: every large product runs on tcgen05) for PNA, GATv2 and TransformerConv.
"""
import copy

import pytest
import torch

import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import models, ops, pna_fn
from gnnepcsaft_b200._lib import lib
from oracle import pyg_restated as O
from tests.util import GAT_CFG_SMALL, PNA_CFG1, branches_from, deg_hist, make_batch

pytestmark = pytest.mark.gpu
DEV = "cuda"
U_TF32 = 2.0 ** -11
C_BOUND = 2.0


def depth_fwd(cfg) -> int:
    """Dependent products on the longest forward path (reference formulation)."""
    L = cfg["propagation_depth"]
    if cfg["conv"] == "PNA":
        per_layer = cfg["pre_layers"] + cfg["post_layers"] + 1          # pre_nn, post_nn, lin
    else:
        per_layer = 2                                                    # node Linears, then the attention sum
    return L * per_layer + 3                                             # + MLP head


def _rel(a, ref, floor=0.0):
    a, ref = a.detach().double().cpu(), ref.detach().double().cpu()
    assert a.shape == ref.shape
    assert bool(torch.isfinite(a).all()), "non-finite values"
    scale = max(float(ref.abs().max()) if ref.numel() else 0.0, floor, 1e-30)
    return (float((a - ref).abs().max()) if a.numel() else 0.0) / scale


def run_tf32_step(cfg, n_graphs, target, seed=1234):
    """One training step on the tf32 tier and on the fp64 oracle (teacher-forced); returns the per-tensor relative
    errors ``{name: (rel_err, bound)}`` and the number of tcgen05 launches."""
    graphs, b = make_batch(n_graphs, seed=seed)
    cfg = dict(cfg, deg=deg_hist(graphs))
    torch.manual_seed(1)
    mo = O.GNNePCSAFT(cfg)
    mp = G.GNNePCSAFT(copy.deepcopy(cfg)).to(DEV)
    mp.load_state_dict(mo.state_dict(), strict=True)
    mo64 = mo.double()
    mo64.train(); mp.train()
    D = depth_fwd(cfg)
    band = C_BOUND * U_TF32 * D
    pna_fn.CAPTURE_AGG, models.CAPTURE_HEAD = [], []
    relu_out = []
    hooks = [bn.register_forward_hook(lambda _m, _i, out: relu_out.append(out.detach())) for bn in mp.batch_norms]
    tc0 = lib().tc_launch_count()
    try:
        pp = mp(b["x"].to(DEV), b["edge_index"].to(DEV), b["edge_attr"].to(DEV), b["batch"].to(DEV))
        aggs, heads = list(pna_fn.CAPTURE_AGG), list(models.CAPTURE_HEAD)
    finally:
        pna_fn.CAPTURE_AGG = models.CAPTURE_HEAD = None
        for h in hooks:
            h.remove()
    with branches_from(aggs, cfg["towers"], relu_out, inner_relus=True, head_outputs=heads, relu_band=band):
        p64 = mo64(b["x"], b["edge_index"], b["edge_attr"], b["batch"])
    l64, _ = O.training_loss(p64, b[target].double())
    l64.backward()
    out2, dpred = ops.relhuber_fwd_bwd(pp.detach().contiguous(), b[target].to(DEV))
    pp.backward(dpred)
    torch.cuda.synchronize()
    tc = lib().tc_launch_count() - tc0
    res = {"pred": (_rel(pp, p64), band), "loss": (_rel(out2[0], l64), band)}
    p64s = dict(mo64.named_parameters())
    gmax = max(float(p.grad.abs().max()) for p in mo64.parameters())
    for n, p in mp.named_parameters():
        assert p.grad is not None, f"no grad for {n}"
        res["d" + n] = (_rel(p.grad, p64s[n].grad, floor=5e-2 * gmax), 2 * band)
    return res, tc


CASES = [
    ("pna_cfg1", PNA_CFG1, 32, "para", False),
    ("gat_cfg1", GAT_CFG_SMALL, 32, "assoc", False),
    ("tconv_cfg1", dict(GAT_CFG_SMALL, conv="Transformer", num_para=3), 32, "para", False),
    ("pna_h256_l6", dict(PNA_CFG1, hidden_dim=256, propagation_depth=6), 256, "para", True),
    ("gat_h256_l6", dict(GAT_CFG_SMALL, hidden_dim=256, propagation_depth=6), 256, "assoc", True),
    ("tconv_h256_l6", dict(GAT_CFG_SMALL, conv="Transformer", num_para=3, hidden_dim=256, propagation_depth=6), 256, "para", True),
    ("pna_h256_l6_towers4", dict(PNA_CFG1, hidden_dim=256, propagation_depth=6, towers=4, pre_layers=1, post_layers=2), 256, "para", True),
]


@pytest.fixture(autouse=True)
def _precision():
    torch.set_float32_matmul_precision("highest")
    ops.set_matmul_precision("tf32")
    torch.manual_seed(0)
    yield
    ops.set_matmul_precision("fp32")


@pytest.mark.parametrize("name,cfg,n_graphs,target,expect_tc", CASES, ids=[c[0] for c in CASES])
def test_tf32_training_step_against_fp64_oracle(name, cfg, n_graphs, target, expect_tc):
    res, tc = run_tf32_step(cfg, n_graphs, target)
    bad = {k: (e, bnd) for k, (e, bnd) in res.items() if not e <= bnd}
    worst = max(res.items(), key=lambda kv: kv[1][0] / kv[1][1])
    assert not bad, (f"{name}: {len(bad)}/{len(res)} tensors outside C*u*D (worst {worst[0]}: "
                     f"{worst[1][0]:.3e} vs {worst[1][1]:.3e}): {sorted(bad.items(), key=lambda kv: -kv[1][0])[:6]}")
    if expect_tc:
        # per layer at least one forward, one dX and one dW tensor-core product, whatever the conv (PNA: ~17)
        need = (7 if cfg["conv"] == "PNA" else 3) * cfg["propagation_depth"]
        assert tc >= need, f"{name}: only {tc} tcgen05 launches (< {need}) -- the tensor path was not taken"


def test_tf32_differs_from_fp32_tier_and_uses_tensor_cores():
    """The two tiers are different code: the tf32 result must NOT be bit-identical with the CUDA-core tier's (a
    silent fallback would be), and the fp32 tier must not launch a single tcgen05 kernel."""
    a = torch.randn(4096, 256, device=DEV)
    w = torch.randn(256, 256, device=DEV)
    t0 = lib().tc_launch_count()
    y32 = ops.gemm(a, w, trans_b=True, precision=ops.PREC_FP32)
    assert lib().tc_launch_count() == t0
    ytc = ops.gemm(a, w, trans_b=True, precision=ops.PREC_TF32)
    assert lib().tc_launch_count() == t0 + 1
    assert not torch.equal(y32, ytc)
    ref = a.double() @ w.double().t()
    e32, etc = float((y32.double() - ref).abs().max()), float((ytc.double() - ref).abs().max())
    assert etc <= 3 * U_TF32 * float((a.double().abs() @ w.double().abs().t()).max())
    assert etc > 10 * e32, "a tf32 product is ~2^13 times coarser than an fp32 one: this one was not rounded"
