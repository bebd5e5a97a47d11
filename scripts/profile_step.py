#!/usr/bin/env python
"""One warm-up + N training steps of a bench workload, for ncu launch lists / single-kernel captures.

    python scripts/profile_step.py [--workload cfg2|cfg3] [--precision tf32|fp32] [--steps 1] [--graphs 4096]
"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gnnepcsaft_b200 import create_model, graph, ops, synth  # noqa: E402
from gnnepcsaft_b200._lib import lib  # noqa: E402
from gnnepcsaft_b200.train import Trainer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--precision", default="tf32")
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--graphs", type=int, default=bench.GRAPHS_PER_GPU)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    ops.set_matmul_precision(args.precision)
    cfg = bench.WORKLOADS[args.workload]
    target = "para" if cfg["num_para"] == 3 else "assoc"
    graphs = bench.make_dataset(0, args.graphs)
    store = graph.DeviceGraphStore(synth.pack(graphs), dev)
    deg = store.calc_deg()
    torch.manual_seed(0)
    model = create_model(dict(cfg), deg).to(dev).train()
    trainer = Trainer(model, target=target)
    ids = np.arange(args.graphs)
    trainer.step(store.collate(ids))
    torch.cuda.synchronize()
    n0 = lib().launch_count()
    import time
    torch.cuda.profiler.start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = trainer.step(store.collate(ids))
    t_host = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    torch.cuda.profiler.stop()
    print(f"host issue time {1e3 * t_host / args.steps:.2f} ms/step, wall {1e3 * t_all / args.steps:.2f} ms/step")
    print(f"steps={args.steps} gnnx_launches_per_step={(lib().launch_count() - n0) // args.steps} loss={float(out[0]):.6f}")


if __name__ == "__main__":
    main()
