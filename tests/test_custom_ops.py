"""The ``torch.library`` custom-op face (gnnepcsaft_b200/torch_ops.py): schemas, fake (shape-inference) kernels and
``torch.export`` tracing -- all without a GPU (tracing runs the fake kernels only).

Reference: ``save_exported_program`` (gnnepcsaft/demo/utils.py:890-915) exports the model with
``torch.export.export(model, (x, edge_index, edge_attr, None), dynamic_shapes={x: (AUTO, 9), edge_index: (2, AUTO),
edge_attr: (AUTO, 3), batch: None})``.
"""
import copy

import pytest
import torch
from torch.export import Dim

import gnnepcsaft_b200 as G
from gnnepcsaft_b200 import torch_ops  # noqa: F401  (registers torch.ops.gnnx.*)
from tests.util import GAT_CFG_SMALL, PNA_CFG1, deg_hist, make_batch

DYN = {"x": (Dim.AUTO, 9), "edge_index": (2, Dim.AUTO), "edge_attr": (Dim.AUTO, 3), "batch": None}
CFGS = [PNA_CFG1, GAT_CFG_SMALL, dict(GAT_CFG_SMALL, conv="Transformer", num_para=3)]


class _Face(torch.nn.Module):
    """forward() insists on CUDA inputs; the CPU trace enters the custom-op face directly (fake kernels only)."""

    def __init__(self, m):
        super().__init__()
        self.m = m

    def forward(self, x, edge_index, edge_attr, batch):
        return self.m.forward_custom_ops(x, edge_index, edge_attr, batch)


def test_every_op_is_registered_with_a_schema():
    want = ["csr_build", "narrow_u8", "edge_codes", "graph_ptr", "embed_sum", "embed_sum_bwd", "linear", "linear_bwd",
            "bn_relu", "bn_relu_bwd", "segment_pool", "segment_pool_bwd", "pna_conv", "gatv2_conv", "transformer_conv"]
    for name in want:
        op = getattr(torch.ops.gnnx, name).default
        assert "gnnx::" + name in str(op._schema)
    s = str(torch.ops.gnnx.pna_conv.default._schema)
    assert "Tensor[] index" in s and "Tensor[] params" in s and "-> Tensor" in s


@pytest.mark.parametrize("cfg", CFGS, ids=[c["conv"] for c in CFGS])
def test_export_traces_with_dynamic_shapes(cfg):
    graphs, b = make_batch(4, seed=3)
    cfg = dict(cfg, deg=deg_hist(graphs))
    torch.manual_seed(0)
    m = G.GNNePCSAFT(copy.deepcopy(cfg)).eval()
    ep = torch.export.export(_Face(m), (b["x"], b["edge_index"], b["edge_attr"], None), dynamic_shapes=DYN)
    used = {str(n.target) for n in ep.graph.nodes if n.op == "call_function" and "gnnx" in str(n.target)}
    conv_op = {"PNA": "gnnx.pna_conv.default", "GATv2": "gnnx.gatv2_conv.default", "Transformer": "gnnx.transformer_conv.default"}
    for need in ("gnnx.csr_build.default", "gnnx.embed_sum.default", "gnnx.bn_relu.default", "gnnx.segment_pool.default",
                 "gnnx.linear.default", conv_op[cfg["conv"]]):
        assert need in used, (need, used)
    n_conv = sum(1 for n in ep.graph.nodes if n.op == "call_function" and str(n.target) == conv_op[cfg["conv"]])
    assert n_conv == cfg["propagation_depth"]
    out = [n for n in ep.graph.nodes if n.op == "output"][0].args[0][0]
    assert tuple(out.meta["val"].shape) == (1, cfg["num_para"])           # batch=None: one graph
    # the node / edge counts stayed symbolic: one exported program serves every molecule
    ph = {n.name: n.meta["val"] for n in ep.graph.nodes if n.op == "placeholder" and n.name in ("x", "edge_index")}
    assert not isinstance(ph["x"].shape[0], int) and not isinstance(ph["edge_index"].shape[1], int)


def test_fake_kernels_infer_shapes():
    from torch._subclasses.fake_tensor import FakeTensorMode
    with FakeTensorMode():
        ei = torch.empty((2, 37), dtype=torch.int64)
        parts = torch.ops.gnnx.csr_build(ei, 11)
        assert [tuple(p.shape) for p in parts] == [(12,), (37,), (37,), (37,), (12,), (37,), (37,), (8,)]
        assert all(p.dtype == torch.int32 for p in parts)
        x = torch.empty((11, 64))
        y = torch.ops.gnnx.bn_relu(x, None, None, torch.empty(64), torch.empty(64), True, 0.1, 1e-5, True, 0.0, 0)
        assert [tuple(t.shape) for t in y] == [(11, 64), (64,), (64,), (64,), (64,)]
        assert tuple(torch.ops.gnnx.linear(x, torch.empty(32, 64), None, True).shape) == (11, 32)
        gp = torch.ops.gnnx.graph_ptr(torch.empty(11, dtype=torch.int64), 3)
        assert tuple(gp.shape) == (4,) and gp.dtype == torch.int32
        assert tuple(torch.ops.gnnx.segment_pool(x, gp, 0).shape) == (3, 64)
