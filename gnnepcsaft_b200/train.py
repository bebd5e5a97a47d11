"""Training step for the hot path: forward, relative-error Huber loss, backward, (one NCCL all-reduce of the flat
gradient), AdamW-amsgrad -- the body of ``GNNePCSAFTL.training_step`` + ``configure_optimizers``
(reference: gnnepcsaft/train/models.py:47-108) without the Lightning shell.

Data parallelism: one process per GPU, graphs sharded by rank, per-rank BatchNorm statistics
(the reference's DDP does not use SyncBN), and exactly ONE all-reduce per step over a single contiguous fp32
gradient buffer; the 1/world_size of the mean is folded into the optimizer kernel.
"""

from __future__ import annotations

import math
from typing import Optional

import torch
import torch.distributed as dist
from torch import Tensor

from . import ops
from .graph import Batch


class FlatParameters:
    """Re-homes every parameter (and its gradient) of ``model`` into one contiguous fp32 buffer each."""

    def __init__(self, model: torch.nn.Module):
        self.params = [p for p in model.parameters() if p.requires_grad]
        if not self.params:
            raise ValueError("model has no trainable parameters")
        dev = self.params[0].device
        sizes = [p.numel() for p in self.params]
        # keep every view 16-byte aligned so vectorised kernels can run on parameter slices
        self.offsets, off = [], 0
        for n in sizes:
            self.offsets.append(off)
            off += (n + 3) // 4 * 4
        self.numel = off
        self.data = torch.zeros(off, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(off, dtype=torch.float32, device=dev)
        for p, o in zip(self.params, self.offsets):
            n = p.numel()
            self.data[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.data[o:o + n].view(p.shape)
        self.grad_views = [self.grad[o:o + p.numel()].view(p.shape) for p, o in zip(self.params, self.offsets)]

    def zero_grad(self) -> None:
        """Drop the per-parameter gradients: with ``.grad`` unset autograd hands each parameter the tensor the backward
        produced (no per-parameter accumulation kernel); ``gather_grads`` then packs them into the flat buffer."""
        for p in self.params:
            p.grad = None

    def gather_grads(self) -> None:
        """Pack every ``p.grad`` into the flat gradient buffer with one multi-tensor copy; afterwards ``p.grad`` IS the
        matching view of the flat buffer.  Parameters the backward did not reach get a zero gradient."""
        src, dst = [], []
        for p, v in zip(self.params, self.grad_views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                src.append(p.grad.reshape(v.shape))
                dst.append(v)
        if src:
            torch._foreach_copy_(dst, src)
        for p, v in zip(self.params, self.grad_views):
            p.grad = v


def shard_range(num_items: int, rank: int, world: int):
    """Contiguous, near-equal shard ``[lo, hi)`` of ``num_items`` for ``rank`` (graphs are independent units:
    sharding needs no data-path collective)."""
    base, rem = divmod(num_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def node_balanced_batches(node_counts, max_nodes: int, max_graphs: int):
    """Greedy split of a graph sequence into consecutive batches of at most ``max_nodes`` nodes / ``max_graphs``
    graphs (bulk screening mixes 1..100-atom molecules, so batches are sized by nodes, not by graph count).
    Returns a list of ``(lo, hi)`` index ranges covering ``range(len(node_counts))`` in order."""
    out, lo, acc = [], 0, 0
    for i, n in enumerate(node_counts):
        n = int(n)
        if i > lo and (acc + n > max_nodes or i - lo >= max_graphs):
            out.append((lo, i))
            lo, acc = i, 0
        acc += n
    if lo < len(node_counts):
        out.append((lo, len(node_counts)))
    return out


def cosine_warm_restarts_lr(base_lr: float, epoch: int, t_0: int, t_mult: int = 2, eta_min: float = 1e-6) -> float:
    """``CosineAnnealingWarmRestarts(T_0=warmup_steps, T_mult=2, eta_min=1e-6)`` as a scalar (models.py:66-74)."""
    t_i, t_cur = t_0, epoch
    while t_cur >= t_i:
        t_cur -= t_i
        t_i *= t_mult
    return eta_min + (base_lr - eta_min) * (1 + math.cos(math.pi * t_cur / t_i)) / 2


class Trainer:
    """One object per rank.  ``step(batch)`` runs a full training step and returns a device tensor
    ``[loss, mape]`` (no host sync inside)."""

    def __init__(self, model: torch.nn.Module, lr: float = 1e-3, weight_decay: float = 1e-2, betas=(0.9, 0.999),
                 eps: float = 1e-5, target: str = "para", huber_delta: float = 0.01,
                 process_group: Optional[dist.ProcessGroup] = None, distributed: Optional[bool] = None):
        self.model = model
        self.flat = FlatParameters(model)
        self.lr, self.weight_decay, self.betas, self.eps = lr, weight_decay, betas, eps
        self.target, self.huber_delta = target, huber_delta
        self.exp_avg = torch.zeros_like(self.flat.data)
        self.exp_avg_sq = torch.zeros_like(self.flat.data)
        self.max_exp_avg_sq = torch.zeros_like(self.flat.data)
        self.step_count = 0
        self.pg = process_group
        self.distributed = dist.is_initialized() if distributed is None else distributed
        self.world = dist.get_world_size(self.pg) if self.distributed else 1
        if self.distributed:                                  # identical replicas to start from
            dist.broadcast(self.flat.data, src=0, group=self.pg)

    def forward_backward(self, batch: Batch) -> Tensor:
        self.flat.zero_grad()
        pred = self.model.forward_batch(batch)
        target = getattr(batch, self.target)
        out2, dpred = ops.relhuber_fwd_bwd(pred.detach().contiguous(), target.contiguous(), self.huber_delta)
        pred.backward(dpred)
        self.flat.gather_grads()
        return out2

    def optimizer_step(self) -> None:
        if self.distributed and self.world > 1:
            dist.all_reduce(self.flat.grad, op=dist.ReduceOp.SUM, group=self.pg)   # the single collective of the step
        self.step_count += 1
        ops.adamw_amsgrad_(self.flat.data, self.flat.grad, self.exp_avg, self.exp_avg_sq, self.max_exp_avg_sq,
                           lr=self.lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                           weight_decay=self.weight_decay, step=self.step_count, grad_scale=1.0 / self.world)

    def step(self, batch: Batch) -> Tensor:
        out2 = self.forward_backward(batch)
        self.optimizer_step()
        return out2
