"""Committed fixtures (tests/golden/*.npz, made by scripts/make_golden.py from the CPU oracle).

CPU leg  : the oracle still reproduces the fixture bit-for-bit-class (rel 1e-6) -- pins the oracle against drift.
GPU leg  : the CUDA path, fed ONLY the fixture (no oracle import at run time), matches its CSR, prediction, loss
           and gradients.  Tolerances: integers bit-exact; prediction / loss rel 1e-4 (fp32, 2 conv layers + head);
           gradients normwise rel 2e-3 per tensor (no teacher forcing here, so a min/max near-tie may route a
           gradient to a different message than the CPU did -- tests/test_gpu_model.py covers those entry by entry).
"""
import json
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["pna_h32_l2", "gatv2_h32_l2"]


def _load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    cfg = json.loads(bytes(z["config"]).decode())
    return z, cfg


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_golden(name):
    from oracle import pyg_restated as O
    z, cfg = _load(name)
    torch.set_float32_matmul_precision("highest")
    torch.set_num_threads(1)
    t = torch.from_numpy
    target = cfg.pop("target")
    model = O.GNNePCSAFT(dict(cfg, deg=z["deg"].tolist())).train()
    model.load_state_dict({k[6:]: t(z[k]) for k in z.files if k.startswith("state/")}, strict=True)
    pred = model(t(z["x"]), t(z["edge_index"]), t(z["edge_attr"]), t(z["batch"]))
    loss, _ = O.training_loss(pred, t(z["target"]))
    loss.backward()
    np.testing.assert_allclose(pred.detach().numpy(), z["pred"], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(loss.item(), z["loss"], rtol=1e-6)
    csr = O.csr_reference(t(z["edge_index"]), z["x"].shape[0])
    assert np.array_equal(csr["rowptr"].numpy(), z["rowptr"]) and np.array_equal(csr["eid"].numpy(), z["eid"])
    for k, p in model.named_parameters():
        g = z["grad/" + k]
        scale = max(float(np.abs(g).max()), 1e-12)
        assert float(np.abs(p.grad.numpy() - g).max()) <= 1e-5 * scale + 1e-9, k
    assert target in ("para", "assoc")


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_path_matches_golden(name):
    import gnnepcsaft_b200 as G
    from gnnepcsaft_b200 import graph, ops
    z, cfg = _load(name)
    cfg.pop("target")
    ops.set_matmul_precision("fp32")
    dev = "cuda"
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    model = G.create_model(dict(cfg), z["deg"].tolist()).to(dev).train()
    model.load_state_dict({k[6:]: t(z[k]) for k in z.files if k.startswith("state/")}, strict=True)
    ei = t(z["edge_index"])
    idx = graph.GraphIndex.build(ei, z["x"].shape[0], check=True)
    assert np.array_equal(idx.rowptr.cpu().numpy(), z["rowptr"].astype(np.int32))
    assert np.array_equal(idx.csr_eid.cpu().numpy(), z["eid"].astype(np.int32))
    pred = model(t(z["x"]), ei, t(z["edge_attr"]), t(z["batch"]))
    np.testing.assert_allclose(pred.detach().cpu().numpy(), z["pred"], rtol=1e-4, atol=1e-5)
    out2, dpred = ops.relhuber_fwd_bwd(pred.detach().contiguous(), t(z["target"]))
    np.testing.assert_allclose(float(out2[0]), z["loss"], rtol=1e-4)
    np.testing.assert_allclose(float(out2[1]), z["mape"], rtol=1e-4)
    pred.backward(dpred)
    gmax = max(float(np.abs(z[k]).max()) for k in z.files if k.startswith("grad/"))
    for k, p in model.named_parameters():
        g = z["grad/" + k].astype(np.float64)
        got = p.grad.detach().double().cpu().numpy()
        denom = max(float(np.linalg.norm(g)), 1e-3 * gmax * np.sqrt(g.size))
        assert float(np.linalg.norm(got - g)) <= 2e-3 * denom, f"grad {k}"
    sd = model.state_dict()
    for k in z.files:
        if k.startswith("after/"):
            np.testing.assert_allclose(sd[k[6:]].cpu().numpy(), z[k], rtol=1e-4, atol=1e-6)
