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


def upper_half_bucket(name: str, depth: int) -> bool:
    """True for the parameters whose gradients are complete FIRST in a backward pass of ``GNNePCSAFT``: the MLP head and
    the upper half of the conv / batch-norm stack.  Their slice of the flat gradient can be all-reduced while the lower
    layers are still running backward."""
    parts = name.split(".")
    if parts[0] == "mlp":
        return True
    if parts[0] in ("convs", "batch_norms") and len(parts) > 1 and parts[1].isdigit():
        return int(parts[1]) >= (depth + 1) // 2
    return False


class FlatParameters:
    """Re-homes every parameter (and its gradient) of ``model`` into one contiguous fp32 buffer each.

    ``early``: optional predicate on the parameter name.  Parameters it selects are laid out behind the others, so that
    their gradients form ONE contiguous slice ``grad[early_lo:]`` (bucket 1) and the rest ``grad[:early_lo]`` (bucket 0);
    ``arm_early_hook(fn)`` then calls ``fn()`` from inside the backward pass as soon as every bucket-1 gradient exists
    and has been packed into the flat buffer."""

    def __init__(self, model: torch.nn.Module, early=None):
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
        if not named:
            raise ValueError("model has no trainable parameters")
        first = [p for n, p in named if not (early and early(n))]
        last = [p for n, p in named if early and early(n)]
        self.params = first + last
        self.n_first = len(first)
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
        self.early_lo = self.offsets[self.n_first] if self.n_first < len(self.params) else self.numel
        self._early_fn = None
        self._early_seen = 0
        self._early_done = False
        self._hooks = []

    # -- bucket 1 ("early") gradients: packed and handed to a callback while the backward is still running
    def arm_early_hook(self, fn) -> None:
        """Install (once) the per-parameter hooks of bucket 1 and set the callback for the NEXT backward pass."""
        if not self._hooks and self.n_first < len(self.params):
            for p in self.params[self.n_first:]:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_early_grad))
        self._early_fn, self._early_seen, self._early_done = fn, 0, False

    def _on_early_grad(self, _p) -> None:
        self._early_seen += 1
        if self._early_fn is not None and not self._early_done and self._early_seen == len(self.params) - self.n_first:
            self._early_done = True
            self._pack(self.n_first, len(self.params))
            self._early_fn()

    def _pack(self, lo: int, hi: int) -> None:
        src, dst = [], []
        for p, v in zip(self.params[lo:hi], self.grad_views[lo:hi]):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                src.append(p.grad.reshape(v.shape))
                dst.append(v)
        if src:
            torch._foreach_copy_(dst, src)
        for p, v in zip(self.params[lo:hi], self.grad_views[lo:hi]):
            p.grad = v

    def zero_grad(self) -> None:
        """Drop the per-parameter gradients: with ``.grad`` unset autograd hands each parameter the tensor the backward
        produced (no per-parameter accumulation kernel); ``gather_grads`` then packs them into the flat buffer."""
        for p in self.params:
            p.grad = None

    def gather_grads(self) -> None:
        """Pack every ``p.grad`` into the flat gradient buffer with one multi-tensor copy; afterwards ``p.grad`` IS the
        matching view of the flat buffer.  Parameters the backward did not reach get a zero gradient."""
        if self._early_done:                       # bucket 1 was packed (and possibly all-reduced) during the backward
            self._pack(0, self.n_first)
        else:
            self._pack(0, len(self.params))
        self._early_fn = None


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


class LossReadback:
    """Per-step device -> host read of ``[loss, mape]`` without stalling the launch queue: step i's values are copied into
    pinned memory asynchronously and handed out when step i + 1 calls ``push`` (the host then waits on an event that
    completed long ago).  A ``loss.item()`` per step drains the queue every step, and the GPU then idles while the host
    issues the ~60 small launches at the head of the next step (measured: 1.3 ms of a 16.2 ms cfg2 step).  The
    reference logs through Lightning's ``self.log`` (models.py:93-106), which does not synchronise per step either."""

    def __init__(self, device, width: int = 2):
        self.slots = [torch.zeros(width, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.events = [torch.cuda.Event(), torch.cuda.Event()]
        self.n = 0
        self.device = device

    def push(self, values: Tensor):
        """Queue the D2H copy of this step's values; returns the PREVIOUS step's values as a list (None at step 0)."""
        k = self.n & 1
        prev = self.last() if self.n > 0 else None
        self.slots[k].copy_(values.detach().reshape(-1)[: self.slots[k].numel()], non_blocking=True)
        self.events[k].record(torch.cuda.current_stream(self.device))
        self.n += 1
        return prev

    def last(self):
        """Values of the most recently pushed step (waits for its copy)."""
        k = (self.n - 1) & 1
        self.events[k].synchronize()
        return self.slots[k].tolist()


class Trainer:
    """One object per rank.  ``step(batch)`` runs a full training step and returns a device tensor
    ``[loss, mape]`` (no host sync inside)."""

    def __init__(self, model: torch.nn.Module, lr: float = 1e-3, weight_decay: float = 1e-2, betas=(0.9, 0.999),
                 eps: float = 1e-5, target: str = "para", huber_delta: float = 0.01,
                 process_group: Optional[dist.ProcessGroup] = None, distributed: Optional[bool] = None,
                 overlap: bool = True):
        self.model = model
        depth = len(getattr(model, "convs", []))
        self.distributed = dist.is_initialized() if distributed is None else distributed
        self.world = dist.get_world_size(process_group) if self.distributed else 1
        # data parallel: the gradients of the head and the upper half of the stack are complete halfway through the
        # backward; their slice of the flat buffer is all-reduced (asynchronously, on NCCL's stream) under the rest of it
        self.overlap = overlap and self.world > 1 and depth >= 2
        self.flat = FlatParameters(model, (lambda n: upper_half_bucket(n, depth)) if self.overlap else None)
        self._early_work = None
        self.lr, self.weight_decay, self.betas, self.eps = lr, weight_decay, betas, eps
        self.target, self.huber_delta = target, huber_delta
        self.exp_avg = torch.zeros_like(self.flat.data)
        self.exp_avg_sq = torch.zeros_like(self.flat.data)
        self.max_exp_avg_sq = torch.zeros_like(self.flat.data)
        self.step_count = 0
        self.pg = process_group
        if self.distributed:                                  # identical replicas to start from
            dist.broadcast(self.flat.data, src=0, group=self.pg)
        if self.world > 1 and torch.cuda.is_available() and self.flat.data.is_cuda:
            # NCCL's kernels share the SMs with the backward from another stream: the kernels with a grid barrier are then
            # launched as cooperative grids (all CTAs co-resident by the driver's guarantee, csrc/common.cuh: launch_coop)
            from ._lib import lib
            lib().call("gnnx_set_gang_launch", 1)

    def backward(self, pred: Tensor, dpred: Tensor) -> None:
        """``pred.backward(dpred)`` with the gradients landing in the flat buffer (and, data parallel, the all-reduce of
        the upper layers' slice started from inside the backward).  Call ``flat.zero_grad()`` before the forward."""
        if self.overlap:
            self.flat.arm_early_hook(self._reduce_early)
        pred.backward(dpred)
        self.flat.gather_grads()

    def forward_backward(self, batch: Batch) -> Tensor:
        self.flat.zero_grad()
        pred = self.model.forward_batch(batch)
        target = getattr(batch, self.target)
        out2, dpred = ops.relhuber_fwd_bwd(pred.detach().contiguous(), target.contiguous(), self.huber_delta)
        self.backward(pred, dpred)
        return out2

    def _reduce_early(self) -> None:
        """Called from inside the backward once bucket 1 is packed: its all-reduce runs under the lower layers' backward."""
        self._early_work = dist.all_reduce(self.flat.grad[self.flat.early_lo:], op=dist.ReduceOp.SUM, group=self.pg, async_op=True)

    def reduce_gradients(self) -> None:
        """Sum the flat gradient over the ranks: ONE all-reduce, or -- with ``overlap`` -- the same bytes in two slices,
        the first of which was started during the backward (``_reduce_early``)."""
        if not (self.distributed and self.world > 1):
            return
        if self._early_work is not None:
            dist.all_reduce(self.flat.grad[:self.flat.early_lo], op=dist.ReduceOp.SUM, group=self.pg)
            self._early_work.wait()
            self._early_work = None
        else:
            dist.all_reduce(self.flat.grad, op=dist.ReduceOp.SUM, group=self.pg)   # the single collective of the step

    def optimizer_step(self) -> None:
        self.reduce_gradients()
        self.step_count += 1
        ops.adamw_amsgrad_(self.flat.data, self.flat.grad, self.exp_avg, self.exp_avg_sq, self.max_exp_avg_sq,
                           lr=self.lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                           weight_decay=self.weight_decay, step=self.step_count, grad_scale=1.0 / self.world)

    def step(self, batch: Batch) -> Tensor:
        out2 = self.forward_backward(batch)
        self.optimizer_step()
        return out2
