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
