"""Full-size (BASELINE configs[1] / [2]: 4096 graphs, hidden 256) property tests and edge cases of the gnnx path.

At these sizes the CPU oracle takes minutes, so parity is checked through size-independent properties of the
reference's block-diagonal batching (PyG ``Batch.from_data_list`` semantics, train/train.py:59-75, models.py:196-227):
bit-identical reruns (atomic-free kernels), graph independence in eval mode (a molecule's prediction does not depend
on which batch it sits in or where), CSR invariants, aggregator checksums, and agreement of the tcgen05 tf32 tier
with the CUDA-core fp32 tier.  The edge cases (zero-edge batches, one-atom molecules, hub rows, a single large
molecule) are compared with the oracle directly.
"""
import copy

import numpy as np
import pytest
import torch

import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import graph, ops, synth
from gnnepcsaft_b200.train import Trainer
from oracle import pyg_restated as O
from tests.util import PNA_CFG1, GAT_CFG_SMALL, assert_as_accurate, assert_close, deg_hist, reproducible

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0) if torch.cuda.is_available() else None

CFG2 = dict(conv="PNA", hidden_dim=256, propagation_depth=6, dropout=0.0, global_pool="add", num_para=3,
            towers=1, pre_layers=2, post_layers=4, heads=2, add_self_loops=True, model="gnn")
CFG3 = dict(CFG2, conv="GATv2", propagation_depth=7, num_para=2)
TIER_PRED_TOL, TIER_GRAD_COS = 3e-2, 0.9


@pytest.fixture(scope="module")
def full():
    graphs = synth.make_graphs(4096, seed=2235)
    store = graph.DeviceGraphStore(synth.pack(graphs), DEV)
    return graphs, store, store.calc_deg()


def _model(cfg, deg, seed=0):
    torch.manual_seed(seed)
    return G.create_model(copy.deepcopy(cfg), deg).to(DEV)


@pytest.fixture()
def tf32():
    ops.set_matmul_precision("tf32")
    yield
    ops.set_matmul_precision("fp32")


def test_fullsize_csr_invariants(full):
    """4096-graph batch: the device CSR is exactly argsort(dst, stable); rowptr = cumsum(bincount(dst)); the src-sorted
    view is the stable argsort of src; the degree histogram is bincount(in-degree) (train/utils.py:37-60)."""
    graphs, store, deg = full
    b = store.collate(np.arange(4096))
    g = b.index
    ei = torch.stack([b._src32.long(), b._dst32.long()])           # the packed collator emits int32 endpoints, no [2,E] int64
    n, e = b.x.shape[0], ei.shape[1]
    assert n == sum(x.num_nodes for x in graphs) and e == sum(x.edge_index.shape[1] for x in graphs)
    order = torch.argsort(ei[1], stable=True)
    assert torch.equal(g.csr_eid.long(), order)
    assert torch.equal(g.csr_src.long(), ei[0][order])
    cnt = torch.bincount(ei[1], minlength=n)
    assert torch.equal(g.rowptr.long(), torch.cat([cnt.new_zeros(1), cnt.cumsum(0)]))
    assert torch.equal(g.in_degree.long(), cnt)
    inv = torch.empty_like(order)
    inv[order] = torch.arange(e, device=order.device)
    assert torch.equal(g.pos_of_eid.long(), inv)
    # src-sorted view: edges grouped by source (stable in edge id), each given as its slot in the dst-sorted array
    by_src = torch.argsort(ei[0], stable=True)
    assert torch.equal(g.csc_pos.long(), inv[by_src])
    assert torch.equal(g.csc_dst.long(), ei[1][by_src])
    cs = torch.bincount(ei[0], minlength=n)
    assert torch.equal(g.colptr.long(), torch.cat([cs.new_zeros(1), cs.cumsum(0)]))
    hist = torch.bincount(cnt).tolist()
    assert deg[:len(hist)] == hist and not any(deg[len(hist):])
    # block-diagonal: every edge stays inside its graph
    gp = b.graph_ptr.long()
    gid = torch.bucketize(ei, gp[1:], right=True)
    assert torch.equal(gid[0], gid[1])


def test_fullsize_aggregate_checksums(full):
    """[E,256] messages of a 4096-graph batch: min <= mean <= max, std >= 0, sum_i deg_i * mean_i = column sums of the
    messages (a checksum of checksums), and bit-identical reruns."""
    _, store, _ = full
    g = store.collate(np.arange(4096)).index
    e, n, f = g.csr_eid.shape[0], g.rowptr.shape[0] - 1, 256
    torch.manual_seed(3)
    m = torch.randn(e, f, device=DEV)
    agg = ops.pna_aggregate_fwd(m, g.rowptr, 1)
    agg2 = ops.pna_aggregate_fwd(m, g.rowptr, 1)
    assert torch.equal(agg, agg2)
    a = agg.view(n, 4, f)
    has = (g.in_degree > 0)[:, None]
    mean, mn, mx, sd = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    assert bool(((mn <= mean + 1e-6) & (mean <= mx + 1e-6) | ~has).all())
    assert bool((sd >= 0).all()) and bool((a[~has[:, 0]] == 0).all())
    lhs = (mean.double() * g.in_degree.double()[:, None]).sum(0)
    rhs = m.double().sum(0)
    assert float((lhs - rhs).abs().max()) <= 1e-5 * float(m.abs().double().sum(0).max())
    # every min / max is attained by a message of the row
    dst = torch.repeat_interleave(torch.arange(n, device=DEV), g.in_degree.long())
    hit_min = torch.zeros(n, f, device=DEV).index_add_(0, dst, (m == mn[dst]).float())
    hit_max = torch.zeros(n, f, device=DEV).index_add_(0, dst, (m == mx[dst]).float())
    assert bool(((hit_min >= 1) | ~has).all()) and bool(((hit_max >= 1) | ~has).all())


@pytest.mark.parametrize("prec,rtol", [("fp32", 2e-5), ("tf32", 4e-3)])      # tf32: 2 of 2331 entries sat at 2.2e-3 on a B200
@pytest.mark.parametrize("cfg", [CFG2, CFG3], ids=["cfg2_pna", "cfg3_gatv2"])
def test_fullsize_graph_independence_and_permutation(full, cfg, prec, rtol):
    """Eval mode (running BatchNorm statistics): the prediction of a molecule must not depend on the batch around it.
    pred(all 4096)[ids] == pred(batch of ids) and a shuffled batch gives the shuffled predictions -- to fp32 rounding on
    the CUDA-core tier (per-destination sums keep their order; only tile positions of the dense products change), and
    to tf32 rounding on the tcgen05 tier (a product whose row count is not tile-aligned, e.g. the MLP head on 777
    graphs, is served by the fp32 kernel, so the two batches round differently there)."""
    _, store, deg = full
    ops.set_matmul_precision(prec)
    try:
        m = _model(cfg, deg).train()
        ids = np.arange(4096)
        with torch.no_grad():
            for lo in (0, 2048):                                      # give the running statistics real values
                m.forward_batch(store.collate(ids[lo:lo + 2048]))
            m.eval()
            y = m.forward_batch(store.collate(ids))
            assert y.shape == (4096, cfg["num_para"]) and bool(torch.isfinite(y).all())
            sub = np.arange(1000, 1777)
            ys = m.forward_batch(store.collate(sub))
            assert_close(ys, y[1000:1777], rtol, 1e-6, "sub-batch vs full batch")
            perm = np.random.default_rng(0).permutation(4096)
            yp = m.forward_batch(store.collate(perm))
            assert_close(yp, y[torch.from_numpy(perm).to(DEV)], rtol, 1e-6, "shuffled batch")
            assert torch.equal(m.forward_batch(store.collate(ids)), y)
    finally:
        ops.set_matmul_precision("fp32")


def test_fullsize_training_step_deterministic_and_tiers_agree(full):
    """One cfg2 training step at the bench size: loss and every gradient bit-identical across reruns (tf32 tier), and
    the tcgen05 tf32 tier agrees with the CUDA-core fp32 tier normwise (10-bit mantissa products, fp32 accumulation:
    ~36 chained products and 8 normalisations deep: predictions within TIER_PRED_TOL normwise).  Gradients of this
    piecewise-linear model (ReLU, min / max routing, the Std mask) respond to a forward perturbation d like sqrt(d) --
    a fraction ~d of the units switches branch -- so the fp32 oracle already deviates from its own fp64 run by 3e-2 here
    and the tf32 tier by ~0.3 (scripts/diag_tiers.py); teacher-forced parity of every gradient is what the small-size
    tests check, this one guards direction and scale (cosine >= TIER_GRAD_COS, norm ratio within 10 %)."""
    _, store, deg = full
    b = store.collate(np.arange(4096))
    res = {}
    for prec in ("tf32", "tf32", "fp32"):
        ops.set_matmul_precision(prec)
        try:
            m = _model(CFG2, deg).train()
            y = m.forward_batch(b)
            out, _ = ops.relhuber_fwd_bwd(y.detach().contiguous(), b.para)
            # a random cotangent: at random initialisation the training loss sits in its L1 region, where d(loss)/d(pred)
            # is nearly the same for every molecule and the BatchNorm backward cancels all but a residual of it -- the
            # fp32 ORACLE's gradients then deviate from its own fp64 run by 3e-2 (scripts/diag_tiers.py), which makes
            # that case a test of the conditioning, not of the products
            torch.manual_seed(17)
            y.backward(torch.randn_like(y))
            torch.cuda.synchronize()
            g = torch.cat([p.grad.flatten() for p in m.parameters()])
            res.setdefault(prec, []).append((y.detach().clone(), g.clone(), float(out[0])))
        finally:
            ops.set_matmul_precision("fp32")
    (y0, g0, l0), (y1, g1, l1) = res["tf32"]
    assert torch.equal(y0, y1) and torch.equal(g0, g1) and l0 == l1
    yf, gf, lf = res["fp32"][0]
    assert bool(torch.isfinite(g0).all())
    ry = float((y0 - yf).norm() / yf.norm())
    rg = float((g0 - gf).norm() / gf.norm())
    cos = float(torch.dot(g0, gf) / (g0.norm() * gf.norm()))
    print(f"tf32 vs fp32 tier at cfg2: predictions {ry:.2e}, gradients {rg:.2e} (normwise), cosine {cos:.4f}, "
          f"loss {l0:.6f} vs {lf:.6f}")
    assert ry <= TIER_PRED_TOL and abs(l0 - lf) <= TIER_PRED_TOL * abs(lf)
    assert cos >= TIER_GRAD_COS and abs(float(g0.norm() / gf.norm()) - 1.0) <= 0.1


def test_fullsize_trainer_loss_decreases(full, tf32):
    """Ten AdamW(amsgrad) steps on one fixed cfg2 batch through train.Trainer (collate + fwd + loss + bwd + step)."""
    _, store, deg = full
    m = _model(CFG2, deg).train()
    tr = Trainer(m)
    b = store.collate(np.arange(4096))
    losses = [float(tr.step(b)[0]) for _ in range(10)]
    assert all(np.isfinite(losses)) and losses[-1] < losses[0]


# ------------------------------------------------------------------------------------------------ edge cases
def _oracle_pair(cfg, graphs, deg=None):
    cfg = dict(cfg, deg=deg if deg is not None else deg_hist(graphs))
    torch.manual_seed(1)
    mo = O.GNNePCSAFT(cfg)
    mp = G.GNNePCSAFT(copy.deepcopy(cfg)).to(DEV)
    mp.load_state_dict(mo.state_dict(), strict=True)
    return mo, mp


@pytest.mark.parametrize("cfg", [PNA_CFG1, GAT_CFG_SMALL, dict(GAT_CFG_SMALL, conv="Transformer")],
                         ids=["pna", "gatv2", "transformer"])
def test_batch_of_single_atoms_has_no_edges(cfg):
    """48 one-atom molecules: E = 0 (ogb_utils.py:137-139 gives [2,0] / [0,3]); forward, loss and every gradient must
    match the oracle (all aggregates are empty: PNA rows are zero, attention rows see only their self loop)."""
    rng = np.random.Generator(np.random.PCG64(5))
    graphs = [synth.make_graph(1, rng) for _ in range(48)]
    ref = synth.make_graphs(32, 11)                                  # degree histogram of a normal training set
    mo, mp = _oracle_pair(cfg, graphs, deg_hist(ref))
    mo64 = copy.deepcopy(mo).double()
    mo.train(); mp.train(); mo64.train()
    x, ei, ea, batch, para, assoc = (torch.from_numpy(a) for a in synth.collate_numpy(graphs))
    assert ei.shape == (2, 0) and ea.shape == (0, 3)
    po = mo(x, ei, ea, batch)
    p64 = mo64(x, ei, ea, batch)
    pp = mp(x.to(DEV), ei.to(DEV), ea.to(DEV), batch.to(DEV))
    # 48 near-identical rows: BatchNorm channels with (almost) no variance amplify rounding by up to 1/sqrt(eps) = 316,
    # so the bar is "as close to the fp64 oracle as the fp32 oracle is" plus a loose elementwise bound
    assert_as_accurate(pp, po, p64, "fwd, E = 0", factor=8.0)
    assert_close(pp, p64, 1e-3, 1e-5, "fwd, E = 0")
    tgt = para if cfg["num_para"] == 3 else assoc
    lo, _ = O.training_loss(po, tgt)
    l64, _ = O.training_loss(p64, tgt.double())
    lo.backward(); l64.backward()
    out, dpred = ops.relhuber_fwd_bwd(pp.detach().contiguous(), tgt.to(DEV))
    assert_close(out[0], l64, 1e-4, 1e-9, "loss, E = 0")
    pp.backward(dpred)
    go, g64 = dict(mo.named_parameters()), dict(mo64.named_parameters())
    for n, p in mp.named_parameters():
        z = torch.zeros_like(go[n])
        r32 = go[n].grad if go[n].grad is not None else z
        r64 = g64[n].grad if g64[n].grad is not None else z.double()
        got = p.grad if p.grad is not None else torch.zeros_like(p)
        assert_as_accurate(got, r32, r64, f"grad {n}, E = 0", factor=8.0, min_scale=1e-6)
        if float(r64.abs().max()) == 0.0:
            assert float(got.abs().max()) <= 1e-6, f"grad {n} must vanish without edges"


def test_one_atom_molecule_and_one_large_molecule_eval():
    """batch=None path (models.py:217-224 pools the whole node set) on a 1-atom and on a 100-atom molecule."""
    rng = np.random.Generator(np.random.PCG64(9))
    ref = synth.make_graphs(32, 11)
    mo, mp = _oracle_pair(PNA_CFG1, ref)
    mo.eval(); mp.eval()
    t = torch.from_numpy
    with torch.no_grad():
        for n in (1, 2, 100):
            g = synth.make_graph(n, rng)
            so = mo(t(g.x), t(g.edge_index), t(g.edge_attr), None)
            sp = mp(t(g.x).to(DEV), t(g.edge_index).to(DEV), t(g.edge_attr).to(DEV), None)
            assert_close(sp, so, 1e-5, 1e-6, f"{n}-atom molecule, batch=None")
            so = mo.pred_with_bounds(O.collate([O.Graph(t(g.x), t(g.edge_index), t(g.edge_attr))]))
            sp = mp.pred_with_bounds(graph.Batch(x=t(g.x).to(DEV), edge_index=t(g.edge_index).to(DEV),
                                                 edge_attr=t(g.edge_attr).to(DEV), batch=torch.zeros(n, dtype=torch.long, device=DEV)))
            assert_close(sp, so, 1e-5, 1e-6, f"{n}-atom molecule, pred_with_bounds")


@pytest.mark.parametrize("F", [64, 256])
def test_aggregate_hub_rows_and_empty_rows(F):
    """Ragged rows: in-degree 0, 1, 2, 31, 32, 33, 100 and 1000 in one call (the reference's scatter reductions have no
    row-length limit); forward against a plain per-row torch evaluation, backward against autograd of it."""
    degs = [0, 1, 2, 0, 31, 32, 33, 100, 0, 1000, 1, 0]
    rowptr = torch.tensor(np.concatenate([[0], np.cumsum(degs)]), dtype=torch.int32, device=DEV)
    e, n = int(rowptr[-1]), len(degs)
    torch.manual_seed(F)
    m = torch.randn(e, F, device=DEV, dtype=torch.float32)
    dst = torch.repeat_interleave(torch.arange(n), torch.tensor(degs))

    def oracle(dtype):
        # the oracle's aggregators (PyG scatter / StdAggregation restated, oracle/pyg_restated.py R3-R4)
        mo = m.detach().cpu().to(dtype).requires_grad_(True)
        out = torch.cat([O._AGGRS[a](mo.view(e, 1, F), dst, n) for a in ("mean", "min", "max", "std")], -1).reshape(n, 4 * F)
        return mo, out
    m32, ref32 = reproducible(lambda: oracle(torch.float32))
    m64, ref64 = oracle(torch.float64)
    agg = ops.pna_aggregate_fwd(m, rowptr, 1)
    assert_close(agg, ref64, 1e-5, 1e-6, "aggregate fwd, ragged rows")
    w = torch.randn(n, 4 * F)
    ref32.backward(w)
    ref64.backward(w.double())
    dm = ops.pna_aggregate_bwd(w.to(DEV), m, agg, rowptr, 1)
    # d std / d m = (m - mean) / (cnt * std) is ill-conditioned where std is small (the two-message row): anchor on fp64
    assert_as_accurate(dm, m32.grad, m64.grad, "aggregate bwd, ragged rows")
    assert_close(dm, m64.grad, 1e-3, 1e-6, "aggregate bwd, ragged rows (elementwise)")
