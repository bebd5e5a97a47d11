"""Host-side data-parallel logic on CPU with the gloo backend, world_size 2.

What is covered here (no GPU): contiguous graph sharding, the flat gradient buffer and its single all-reduce
(sum over ranks == gradient of the global batch; the 1/world scale is folded into the optimizer kernel on the
GPU), the dataset degree histogram as a sum of per-shard histograms, and node-balanced inference batches.
"""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gnnepcsaft_b200 import synth
from gnnepcsaft_b200.train import FlatParameters, node_balanced_batches, shard_range


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        model = torch.nn.Sequential(torch.nn.Linear(9, 16), torch.nn.ReLU(), torch.nn.Linear(16, 3))
        flat = FlatParameters(model)
        dist.broadcast(flat.data, src=0)
        graphs = synth.make_graphs(37, seed=99)                       # 37 graphs: uneven shards (19 / 18)
        feats = torch.stack([torch.from_numpy(g.x).float().mean(0) for g in graphs])
        tgt = torch.stack([torch.from_numpy(g.para[0]) for g in graphs])
        lo, hi = shard_range(len(graphs), rank, world)
        flat.zero_grad()
        ((model(feats[lo:hi]) - tgt[lo:hi]) ** 2).sum().backward()
        flat.gather_grads()
        assert all(p.grad.data_ptr() == flat.grad.data_ptr() + 4 * o for p, o in zip(flat.params, flat.offsets)), \
            "gradients must land in the flat buffer (one all-reduce, no bucketing)"
        dist.all_reduce(flat.grad, op=dist.ReduceOp.SUM)              # THE collective of a training step
        # degree histogram of the dataset = sum of shard histograms
        hist = torch.zeros(16, dtype=torch.int64)
        for g in graphs[lo:hi]:
            deg = np.bincount(g.edge_index[1], minlength=g.num_nodes)
            hist += torch.from_numpy(np.bincount(deg, minlength=16)[:16])
        dist.all_reduce(hist)
        if rank == 0:
            torch.save(dict(grad=flat.grad.clone(), hist=hist, shards=[shard_range(len(graphs), r, world) for r in range(world)]), out)
    finally:
        dist.destroy_process_group()


def test_two_rank_sharded_gradient_equals_global_batch(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    assert got["shards"] == [(0, 19), (19, 37)]
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(9, 16), torch.nn.ReLU(), torch.nn.Linear(16, 3))
    flat = FlatParameters(model)
    graphs = synth.make_graphs(37, seed=99)
    feats = torch.stack([torch.from_numpy(g.x).float().mean(0) for g in graphs])
    tgt = torch.stack([torch.from_numpy(g.para[0]) for g in graphs])
    flat.zero_grad()
    ((model(feats) - tgt) ** 2).sum().backward()
    flat.gather_grads()
    assert torch.allclose(got["grad"], flat.grad, rtol=1e-5, atol=1e-3 * float(flat.grad.abs().max()) * 1e-2)
    hist = np.zeros(16, np.int64)
    for g in graphs:
        hist += np.bincount(np.bincount(g.edge_index[1], minlength=g.num_nodes), minlength=16)[:16]
    assert got["hist"].tolist() == hist.tolist()


def test_shard_range_covers_everything_once():
    for n in (0, 1, 7, 4096, 10_000_001):
        for w in (1, 2, 3, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1


def test_node_balanced_batches():
    rng = np.random.default_rng(0)
    n = rng.integers(1, 101, size=5000)
    batches = node_balanced_batches(n, max_nodes=4096, max_graphs=128)
    assert batches[0][0] == 0 and batches[-1][1] == len(n)
    assert all(a[1] == b[0] for a, b in zip(batches, batches[1:]))
    for lo, hi in batches:
        assert hi - lo <= 128 and (n[lo:hi].sum() <= 4096 or hi - lo == 1)
    assert node_balanced_batches([], 10, 10) == []
    assert node_balanced_batches([500], 10, 10) == [(0, 1)]


class _Toy(torch.nn.Module):
    """Names like GNNePCSAFT's: convs.{l}.*, batch_norms.{l}.*, mlp.*"""

    def __init__(self):
        super().__init__()
        self.node_embed = torch.nn.Linear(9, 16)
        self.convs = torch.nn.ModuleList([torch.nn.Linear(16, 16) for _ in range(4)])
        self.batch_norms = torch.nn.ModuleList([torch.nn.LayerNorm(16) for _ in range(4)])
        self.mlp = torch.nn.Linear(16, 3)

    def forward(self, x):
        h = self.node_embed(x)
        for c, b in zip(self.convs, self.batch_norms):
            h = torch.relu(b(c(h)))
        return self.mlp(h)


def _overlap_worker(rank, world, port, out):
    from gnnepcsaft_b200.train import upper_half_bucket
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        res = {}
        x, y = torch.randn(40, 9), torch.randn(40, 3)
        lo, hi = shard_range(40, rank, world)
        for mode in ("single", "two_buckets"):
            torch.manual_seed(1)
            model = _Toy()
            flat = FlatParameters(model, (lambda n: upper_half_bucket(n, 4)) if mode == "two_buckets" else None)
            dist.broadcast(flat.data, src=0)
            fired = {}

            def early():
                # called from INSIDE the backward: the lower layers have no gradient yet, the upper slice is packed
                fired["lower_pending"] = model.convs[0].weight.grad is None
                fired["work"] = dist.all_reduce(flat.grad[flat.early_lo:], async_op=True)

            for step in range(2):                              # twice: the hooks persist, the callback is re-armed
                flat.zero_grad()
                fired.clear()
                if mode == "two_buckets":
                    flat.arm_early_hook(early)
                ((model(x[lo:hi]) - y[lo:hi]) ** 2).sum().backward()
                flat.gather_grads()
                if mode == "two_buckets":
                    assert fired.get("lower_pending") is True, "the early slice must be reduced while the backward is still running"
                    dist.all_reduce(flat.grad[:flat.early_lo])
                    fired["work"].wait()
                else:
                    dist.all_reduce(flat.grad)
            names = {id(p): n for n, p in model.named_parameters()}
            res[mode] = {names[id(p)]: p.grad.clone() for p in flat.params}
            if mode == "two_buckets":
                early_names = [names[id(p)] for p in flat.params[flat.n_first:]]
                assert all(upper_half_bucket(n, 4) for n in early_names) and len(early_names) == 2 * 2 + 2 * 2 + 2
                assert flat.early_lo == flat.offsets[flat.n_first]
        if rank == 0:
            torch.save(res, out)
    finally:
        dist.destroy_process_group()


def test_two_bucket_overlapped_allreduce_equals_the_single_one(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_overlap_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    res = torch.load(out)
    assert set(res["single"]) == set(res["two_buckets"])
    for n, g in res["single"].items():
        assert torch.equal(g, res["two_buckets"][n]), n
