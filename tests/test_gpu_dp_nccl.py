"""Data parallelism on hardware (reference: train/train.py:85-99 -- Lightning DDP over graph
mini-batches, per-rank BatchNorm): two ranks over NCCL, each running ``Trainer.forward_backward`` of the real
``GNNePCSAFT`` on its shard of a global batch; the all-reduced flat gradient times 1/world must equal the gradient of
the global batch on one GPU.  BatchNorm is frozen (eval statistics), as the statement requires: batch statistics are
per rank by design.  Skipped with fewer than two GPUs (run it with ``gpurun --gpus 2``)."""
import copy
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _setup(conv, device):
    import gnnepcsaft_b200 as G
    from gnnepcsaft_b200 import graph, ops, synth
    from tests.util import GAT_CFG_SMALL, PNA_CFG1
    torch.set_float32_matmul_precision("highest")
    ops.set_matmul_precision("fp32")
    graphs = synth.make_graphs(32, 4242, 10, 40, 0.0)
    store = graph.DeviceGraphStore(synth.pack(graphs), device)
    cfg = dict(PNA_CFG1 if conv == "PNA" else dict(GAT_CFG_SMALL, num_para=3))
    torch.manual_seed(11)
    model = G.create_model(copy.deepcopy(cfg), store.calc_deg()).to(device)
    # BatchNorm frozen at non-trivial statistics; everything else in training mode
    for m in model.modules():
        if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
            m.running_mean.uniform_(-0.1, 0.1)
            m.running_var.uniform_(0.8, 1.2)
            m.eval()
    return model, store


def _worker(rank, world, port, conv, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from gnnepcsaft_b200.train import Trainer, shard_range
        model, store = _setup(conv, dev)
        tr = Trainer(model)                                        # distributed: broadcasts rank 0's parameters
        assert tr.world == world
        lo, hi = shard_range(store.num_graphs, rank, world)
        tr.forward_backward(store.collate(np.arange(lo, hi)))
        assert tr.overlap and tr._early_work is not None, "the upper layers' slice must already be in flight after the backward"
        tr.reduce_gradients()                                      # the rest of the step's collective (Trainer.optimizer_step)
        torch.cuda.synchronize()
        if rank == 0:
            # by NAME: the two-bucket trainer lays its flat buffer out differently from the single-GPU one
            sharded = {n: (p.grad / world).cpu() for n, p in model.named_parameters()}
            model1, store1 = _setup(conv, dev)
            model1.load_state_dict(model.state_dict())
            one = Trainer(model1, distributed=False)
            one.forward_backward(store1.collate(np.arange(store1.num_graphs)))
            torch.cuda.synchronize()
            torch.save(dict(sharded=sharded, whole={n: p.grad.cpu() for n, p in model1.named_parameters()}), out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("conv", ["PNA", "GATv2"])
def test_two_rank_nccl_gradient_equals_global_batch(conv, tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    out = str(tmp_path / "grad.pt")
    mp.spawn(_worker, args=(2, _free_port(), conv, out), nprocs=2, join=True)
    got = torch.load(out)
    assert set(got["sharded"]) == set(got["whole"])
    gmax = max(float(b.abs().max()) for b in got["whole"].values())
    assert gmax > 0
    # equal shards (16 + 16 graphs): mean of the shard means == global mean; fp32 summation order differs
    for n, b in got["whole"].items():
        a = got["sharded"][n]
        assert float((a.double() - b.double()).abs().max()) <= 1e-4 * gmax, (n, float((a - b).abs().max()) / gmax)
