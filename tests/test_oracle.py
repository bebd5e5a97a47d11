"""CPU tests that pin the oracle itself: hand-computed cases and fp64 gradchecks."""
import math

import torch

from oracle import pyg_restated as O


def test_scatter_minmax_empty_and_ties():
    src = torch.tensor([[1.0], [1.0], [0.5], [-2.0]], requires_grad=True)
    idx = torch.tensor([0, 0, 0, 2])
    out = O.scatter(src, idx, 4, "max")
    assert out.flatten().tolist() == [1.0, 0.0, -2.0, 0.0]          # empty rows -> 0
    out.sum().backward()
    assert src.grad.flatten().tolist() == [0.5, 0.5, 0.0, 1.0]      # even split among ties
    assert O.scatter(src.detach(), idx, 4, "min").flatten().tolist() == [0.5, 0.0, -2.0, 0.0]
    assert torch.allclose(O.scatter(src.detach(), idx, 4, "mean").flatten(), torch.tensor([2.5 / 3, 0.0, -2.0, 0.0]))


def test_std_threshold():
    # one message per node -> var 0 -> std exactly 0 with zero gradient
    x = torch.tensor([[3.0], [1.0], [2.0]], requires_grad=True)
    out = O.std_aggregation(x, torch.tensor([0, 1, 1]), 2)
    assert out[0].item() == 0.0
    assert abs(out[1].item() - 0.5) < 1e-7
    out.sum().backward()
    assert x.grad[0].item() == 0.0
    assert abs(x.grad[1].item() + 0.5) < 1e-6 and abs(x.grad[2].item() - 0.5) < 1e-6


def test_segment_softmax_hand():
    a = torch.tensor([0.0, math.log(3.0), 5.0])
    out = O.segment_softmax(a, torch.tensor([0, 0, 1]), 3)
    assert torch.allclose(out, torch.tensor([0.25, 0.75, 1.0]), atol=1e-7)


def test_degree_scaler_three_nodes():
    deg = torch.tensor([0, 2, 1])            # two nodes of degree 1, one of degree 2
    mod = O.DegreeScalerAggregation(["mean", "min", "max", "std"], ["identity", "amplification", "attenuation"], deg)
    delta = (2 * math.log(2) + math.log(3)) / 3
    assert abs(mod.avg_deg_log.item() - delta) < 1e-6
    m = torch.tensor([[1.0], [3.0], [5.0]])
    out = mod(m, torch.tensor([0, 0, 1]), 3)   # node0 <- {1,3}, node1 <- {5}, node2 <- {}
    ident = torch.tensor([[2.0, 1.0, 3.0, 1.0], [5.0, 5.0, 5.0, 0.0], [0.0, 0.0, 0.0, 0.0]])
    assert torch.allclose(out[:, :4], ident, atol=1e-6)
    assert torch.allclose(out[0, 4:8], ident[0] * math.log(3) / delta, atol=1e-5)
    assert torch.allclose(out[1, 8:12], ident[1] * delta / math.log(2), atol=1e-5)
    assert out[2].abs().sum().item() == 0.0


def test_collate_and_csr():
    g0 = O.Graph(torch.zeros(3, 9, dtype=torch.long), torch.tensor([[0, 1, 1, 2], [1, 0, 2, 1]]), torch.zeros(4, 3, dtype=torch.long))
    g1 = O.Graph(torch.zeros(1, 9, dtype=torch.long), torch.zeros(2, 0, dtype=torch.long), torch.zeros(0, 3, dtype=torch.long))
    g2 = O.Graph(torch.zeros(2, 9, dtype=torch.long), torch.tensor([[0, 1], [1, 0]]), torch.zeros(2, 3, dtype=torch.long))
    b = O.collate([g0, g1, g2])
    assert b.edge_index.tolist() == [[0, 1, 1, 2, 4, 5], [1, 0, 2, 1, 5, 4]]
    assert b.batch.tolist() == [0, 0, 0, 1, 2, 2] and b.ptr.tolist() == [0, 3, 4, 6]
    csr = O.csr_reference(b.edge_index, 6)
    assert csr["rowptr"].tolist() == [0, 1, 3, 4, 4, 5, 6]
    assert csr["eid"].tolist() == [1, 0, 3, 2, 5, 4]
    assert O.calc_deg([g0, g1, g2]) == [1, 4, 1]


def _gc(fn, *inputs):
    return torch.autograd.gradcheck(fn, inputs, eps=1e-6, atol=1e-5, rtol=1e-4)


def test_gradcheck_pna_fp64():
    torch.manual_seed(0)
    conv = O.PNAConv(8, 8, ["mean", "min", "max", "std"], ["identity", "amplification", "attenuation"],
                     torch.tensor([0, 3, 2, 1]), edge_dim=8, towers=2, pre_layers=2, post_layers=2, divide_input=True).double()
    ei = torch.tensor([[0, 1, 1, 2, 3, 1, 0, 3], [1, 0, 2, 1, 1, 3, 3, 0]])
    x = torch.randn(5, 8, dtype=torch.double, requires_grad=True)
    ea = torch.randn(8, 8, dtype=torch.double, requires_grad=True)
    assert _gc(lambda a, b: conv(a, ei, b), x, ea)


def test_gradcheck_gatv2_fp64():
    torch.manual_seed(0)
    conv = O.GATv2Conv(8, 4, heads=2, edge_dim=8, add_self_loops=True).double()
    ei = torch.tensor([[0, 1, 1, 2, 3, 1], [1, 0, 2, 1, 1, 3]])
    x = torch.randn(5, 8, dtype=torch.double, requires_grad=True)
    ea = torch.randn(6, 8, dtype=torch.double, requires_grad=True)
    assert _gc(lambda a, b: conv(a, ei, b), x, ea)


def test_gatv2_loop_attr_is_mean_of_incoming():
    ei = torch.tensor([[0, 1, 2, 2], [1, 0, 1, 2]])       # last edge is a self loop and gets removed
    ea = torch.tensor([[1.0], [2.0], [4.0], [9.0]])
    ei2, ea2 = O.remove_self_loops(ei, ea)
    ei3, ea3 = O.add_self_loops_mean(ei2, ea2, 3)
    assert ei3.tolist() == [[0, 1, 2, 0, 1, 2], [1, 0, 1, 0, 1, 2]]
    assert ea3.flatten().tolist() == [1.0, 2.0, 4.0, 2.0, 2.5, 0.0]


def test_model_shapes_and_bounds():
    torch.manual_seed(0)
    cfg = dict(conv="PNA", hidden_dim=16, propagation_depth=2, dropout=0.0, global_pool="add", num_para=3,
               towers=2, pre_layers=1, post_layers=1, deg=[0, 2, 1])
    m = O.GNNePCSAFT(cfg).eval()
    g = O.Graph(torch.zeros(3, 9, dtype=torch.long), torch.tensor([[0, 1, 1, 2], [1, 0, 2, 1]]), torch.zeros(4, 3, dtype=torch.long))
    out = m.pred_with_bounds(g)                       # batch=None single-graph path
    assert out.shape == (1, 3)
    assert bool((out >= m.lower_bounds[:3]).all()) and bool((out <= m.upper_bounds[:3]).all())
