"""Which CPU formulation of the oracle's segment sums is reproducible on the GPU box's host (16 threads)?
tests/test_gpu_ops.py::test_pna_aggregate_fwd_bwd failed intermittently because the fp32 ORACLE's std aggregate moved by
5e-4 between two evaluations of the same input (the CUDA result matched the fp64 oracle both times)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tests.util import make_batch

torch.manual_seed(0)
_, b = make_batch(48, seed=9)
idx, N = b["edge_index"][1], b["x"].shape[0]
E, F = idx.shape[0], 64
m = torch.randn(E, 1, F)
bi = idx.view(-1, 1, 1).expand_as(m)
ref = torch.zeros(N, 1, F, dtype=torch.float64).index_add_(0, idx, (m * m).double())
forms = {
    "scatter_add_": lambda s: s.new_zeros(N, 1, F).scatter_add_(0, bi, s),
    "scatter_add_ contiguous index": lambda s: s.new_zeros(N, 1, F).scatter_add_(0, bi.contiguous(), s),
    "index_add_": lambda s: s.new_zeros(N, 1, F).index_add_(0, idx, s),
}
print("threads", torch.get_num_threads())
for name, fn in forms.items():
    bad, worst = 0, 0.0
    t0 = time.time()
    for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 400):
        out = fn(m * m)
        err = float((out.double() - ref).abs().max())
        if err > 1e-4:
            bad += 1
        worst = max(worst, err)
    print(f"{name:32s} bad {bad:4d}  worst err {worst:.3e}  {time.time() - t0:.1f}s")
