"""Experiment: k_pna_gather_fwd with neighbour rows staged through shared memory by cp.async.bulk (GNNX_GATHER_BULK=1)
against the direct float4-load kernel, at cfg2's shape: bit-identical output, then time per launch."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gnnepcsaft_b200 import graph, ops, synth  # noqa: E402

DEV = "cuda"
graphs = synth.make_graphs(4096, 7, 10, 40, 0.0)
store = graph.DeviceGraphStore(synth.pack(graphs), torch.device(DEV))
bt = store.collate(np.arange(4096))
g = bt.index
N, E, H = g.num_nodes, g.num_edges, 256
pq = torch.randn(N, 2 * H, device=DEV)
r = torch.randn(60, H, device=DEV)
codes = torch.randint(0, 60, (E,), device=DEV).to(torch.uint8)
res = {}
for mode in ("0", "1"):
    os.environ["GNNX_GATHER_BULK"] = mode
    h, bits = ops.pna_gather_fwd(pq, r, codes, g.rowptr, g.csr_src, relu=True, want_relu_bits=True)
    torch.cuda.synchronize()
    for _ in range(3):
        ops.pna_gather_fwd(pq, r, codes, g.rowptr, g.csr_src, relu=True, want_relu_bits=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        ops.pna_gather_fwd(pq, r, codes, g.rowptr, g.csr_src, relu=True, want_relu_bits=True)
    e1.record()
    torch.cuda.synchronize()
    res[mode] = (h, bits, e0.elapsed_time(e1) * 1e3 / 20)
print(f"N {N} E {E}: direct {res['0'][2]:.1f} us, bulk-staged {res['1'][2]:.1f} us; "
      f"h identical {torch.equal(res['0'][0], res['1'][0])}, bits identical {torch.equal(res['0'][1], res['1'][1])}")
