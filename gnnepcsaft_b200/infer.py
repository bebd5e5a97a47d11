"""Bulk screening: ``pred_with_bounds`` over a device-resident molecule store, sharded with no communication.

The reference predicts one molecule per call (``predict_params_from_inchi`` -> ``_predict_with_model`` ->
``model.pred_with_bounds(graphs)``, gnnepcsaft/demo/utils.py:549-562, 631-649); BASELINE config 5 (10 M molecules
over 8 GPUs) is the same call -- eval mode, ``torch.no_grad()``, clipped to the parameter bounds of
``train/models.py:167-172, 246-253`` -- run over batches collated on the device.  Graphs are independent units, so
each rank screens a contiguous shard (``train.shard_range``) and writes its own slice: zero collectives.
Batches are sized by NODE count (1..100-atom molecules), not by graph count.
"""

from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np
import torch
from torch import Tensor

from . import ops
from .graph import DeviceGraphStore
from .train import node_balanced_batches, shard_range


class BulkScreener:
    """``BulkScreener(model, store).run()`` -> ``[G_shard, num_para]`` clipped predictions of this rank's shard."""

    def __init__(self, model: torch.nn.Module, store: DeviceGraphStore, rank: int = 0, world: int = 1,
                 max_nodes: int = 1 << 20, max_graphs: int = 16384):
        self.model, self.store = model, store
        self.lo, self.hi = shard_range(store.num_graphs, rank, world)
        counts = np.diff(store.node_ptr_host[self.lo:self.hi + 1])
        self.batches: List[Tuple[int, int]] = [(self.lo + a, self.lo + b)
                                               for a, b in node_balanced_batches(counts, max_nodes, max_graphs)]

    @property
    def num_graphs(self) -> int:
        return self.hi - self.lo

    @torch.no_grad()
    def run_batch(self, k: int, out: Optional[Tensor] = None) -> Tensor:
        lo, hi = self.batches[k]
        pred = self.model.pred_with_bounds(self.store.collate(np.arange(lo, hi)))
        if out is not None:
            out[lo - self.lo:hi - self.lo].copy_(pred)
        return pred

    @torch.no_grad()
    def run(self, out: Optional[Tensor] = None) -> Tensor:
        was_training = self.model.training
        self.model.eval()
        try:
            if out is None:
                out = torch.empty((self.num_graphs, self.model.num_para), dtype=torch.float32, device=self.store.device)
            for k in range(len(self.batches)):
                self.run_batch(k, out)
        finally:
            self.model.train(was_training)
        return out


class HostResultStream:
    """Predictions of successive batches streamed into ONE pinned host array without a synchronisation per batch: each
    batch's rows are copied device -> host on a side stream behind an event, and the device tensor is kept alive until
    its copy has drained (``record_stream``).  ``finish()`` waits once.  Replaces ``out.cpu().numpy()`` per shard."""

    def __init__(self, rows: int, cols: int, device):
        self.host = torch.empty((rows, cols), dtype=torch.float32).pin_memory()
        self.stream = torch.cuda.Stream(device)
        self.device = device
        self.bytes = 0

    def push(self, row0: int, pred: Tensor) -> None:
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            self.host[row0:row0 + pred.shape[0]].copy_(pred, non_blocking=True)
        pred.record_stream(self.stream)
        self.bytes += pred.numel() * 4

    def finish(self) -> Tensor:
        self.stream.synchronize()
        return self.host


@torch.no_grad()
def screen_to_host(model: torch.nn.Module, store: DeviceGraphStore, rank: int = 0, world: int = 1,
                   max_nodes: int = 1 << 21, max_graphs: int = 16384) -> Tuple[Tensor, int]:
    """A rank's whole shard through ``pred_with_bounds`` with the results landing in pinned host memory while the next
    batches run -> ``(pinned [G_shard, num_para], bytes copied device -> host)``."""
    scr = BulkScreener(model, store, rank, world, max_nodes, max_graphs)
    was_training = model.training
    model.eval()
    try:
        sink = HostResultStream(scr.num_graphs, model.num_para, store.device)
        for k, (lo, hi) in enumerate(scr.batches):
            sink.push(lo - scr.lo, scr.run_batch(k))
        return sink.finish(), sink.bytes
    finally:
        model.train(was_training)


def screen_parameters(model_msigmae: torch.nn.Module, store: DeviceGraphStore, munanb: Tensor, mw: Tensor,
                      model_assoc: Optional[torch.nn.Module] = None, assoc: Optional[Tensor] = None, rank: int = 0,
                      world: int = 1, max_nodes: int = 1 << 20, max_graphs: int = 16384,
                      out_path: Optional[str] = None) -> Tensor:
    """Bulk-screening output stage (reference ``_get_model_params``, demo/utils.py:664-691, run
    over a whole shard): both models' ``pred_with_bounds`` over this rank's graphs, then ONE kernel assembles the
    9-vector ``[m, sigma, epsilon/k, kappa_AB, epsilon_AB/k, mu, n_a, n_b, MW]`` per molecule --
    ``10 ** (pred * [-1, +1])`` for the association pair when ``model_assoc`` is given, the known ``assoc`` values (or
    zeros) otherwise.  ``munanb`` [G,3] / ``mw`` [G] are the per-graph attributes of the WHOLE store (this rank's slice
    is taken here).  With ``out_path`` the shard is also streamed to ``<out_path>.rank<r>.npy`` (one file per rank:
    no communication)."""
    if model_assoc is not None and assoc is not None:
        raise ValueError("screen_parameters: give `model_assoc` or known `assoc` values, not both")
    scr = BulkScreener(model_msigmae, store, rank, world, max_nodes, max_graphs)
    msigmae = scr.run()
    lo, hi = scr.lo, scr.hi
    log10assoc = None
    if model_assoc is not None:
        log10assoc = BulkScreener(model_assoc, store, rank, world, max_nodes, max_graphs).run()
    dev = store.device
    out = ops.assemble_params(msigmae, munanb[lo:hi].to(dev, torch.float32).contiguous(),
                              mw.reshape(-1)[lo:hi].to(dev, torch.float32).contiguous(), log10assoc=log10assoc,
                              assoc=None if assoc is None else assoc[lo:hi].to(dev, torch.float32).contiguous())
    if out_path is not None:
        np.save(f"{out_path}.rank{rank}.npy", out.cpu().numpy())
    return out
