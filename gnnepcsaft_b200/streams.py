"""A second CUDA stream per device for parameter-only ("weight-space") work queued ahead of the layers that use it.

GNNePCSAFT can compute the parameter-only tensors of all its PNAConv layers (fused ``W_lin W_last`` weight, the
60-row edge-term tables, the per-degree weights: ~5 tiny kernels per layer) before the layer loop.  In eval mode
the result is cached until a parameter changes (bulk screening), so the stream only matters for the first batch; in
training the work is replayed from ONE CUDA graph (``PREP_GRAPH``).  Measured on B200 (cfg2) with eager side-stream
launches: 0.2-0.35 ms off the device-resident step, but +0.7 ms on a step that blocks on its loss, because ~50 launches
are front-loaded while the GPU waits for the host; the graph makes the front-loading one copy and one launch.
(The backward's weight-space chains were also tried on this stream: no gain -- a persistent GEMM CTA fills the SM's
register file, so small kernels cannot run underneath it.)  Ordering is by events only: ``begin`` makes the side
stream wait for everything already queued on the caller's stream, an event from ``mark`` makes the caller's stream
wait for the side work, so results are bit-identical with ``ENABLED = False`` (tests compare both).

Allocator note: tensors created inside a region belong to the side stream's pool.  They are consumed by kernels queued
on the main stream after an event wait, and the next region always starts with ``begin`` (side waits for main), so a
freed block can never be handed to a side kernel that runs before its last main-stream reader.
"""
from __future__ import annotations

import torch

import os

ENABLED = os.environ.get("GNNX_SIDE_STREAM", "1") != "0"
# Training forward: queue every layer's parameter-only work ahead of the layer loop (GNNX_PREP_AHEAD=0: inline, per layer).
PREP_AHEAD = os.environ.get("GNNX_PREP_AHEAD", "1") == "1"
# ... and replay that work from ONE CUDA graph (GNNePCSAFT._prepare_convs_graph): the front-loading is then a copy and a
# graph launch, so the blocking variants of the step no longer pay for it.  GNNX_PREP_GRAPH=0: eager side-stream launches.
PREP_GRAPH = os.environ.get("GNNX_PREP_GRAPH", "1") != "0"
_SIDES: dict = {}


class SideStream:
    def __init__(self, device: torch.device):
        self.stream = torch.cuda.Stream(device)
        self._main = None

    def begin(self) -> None:
        """Fork: the side stream waits for the current stream's tail and becomes torch's current stream."""
        self._main = torch.cuda.current_stream()
        self.stream.wait_stream(self._main)
        torch.cuda.set_stream(self.stream)

    def mark(self) -> torch.cuda.Event:
        ev = torch.cuda.Event()
        ev.record(self.stream)
        return ev

    def end(self) -> None:
        """Back to the caller's stream (no synchronisation: use ``join`` or an event from ``mark``)."""
        torch.cuda.set_stream(self._main)

    def join(self) -> None:
        torch.cuda.current_stream().wait_stream(self.stream)


def side(device: torch.device):
    """The side stream of ``device``, or ``None`` when disabled."""
    if not ENABLED:
        return None
    key = device.index if device.index is not None else torch.cuda.current_device()
    s = _SIDES.get(key)
    if s is None:
        s = _SIDES[key] = SideStream(torch.device("cuda", key))
    return s
