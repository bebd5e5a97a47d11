#!/usr/bin/env python
"""Times the gnnx GEMM layouts at the hot path's shapes (CUDA events, L2 flushed by size: operands >> 126 MB)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gnnepcsaft_b200 import ops  # noqa: E402


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def single(M, N, K, layout):
    """One shape, 3 launches: the target of an `ncu --set full` capture."""
    dev = "cuda"
    ops.set_matmul_precision("tf32")
    a = torch.randn(M, K, device=dev)
    w = torch.randn(N, K, device=dev)
    wt = w.t().contiguous()
    c = torch.empty(M, N, device=dev)
    dy = torch.randn(M, N, device=dev)
    for _ in range(3):
        if layout == "NT":
            ops.gemm(a, w, trans_b=True, out=c)
        elif layout == "NN":
            ops.gemm(a, wt, out=c)
        else:
            ops.gemm(dy, a, trans_a=True)
    torch.cuda.synchronize()
    print("ok")


def main():
    dev = "cuda"
    if len(sys.argv) > 1 and sys.argv[1] == "--single":
        return single(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5])
    ops.set_matmul_precision(sys.argv[1] if len(sys.argv) > 1 else "tf32")
    rows = [102400, 204800]
    shapes = [(256, 256), (256, 1024), (768, 1024), (1024, 256), (512, 256)]     # (N, K)
    print(f"{'layout':8s} {'M':>7s} {'N':>5s} {'K':>5s} {'ms':>8s} {'TF/s':>7s} {'GB/s':>7s}")
    for M in rows:
        for N, K in shapes:
            a = torch.randn(M, K, device=dev)
            w = torch.randn(N, K, device=dev)
            wt = w.t().contiguous()
            c = torch.empty(M, N, device=dev)
            flops = 2.0 * M * N * K
            byts = 4.0 * (M * K + M * N + N * K)
            for name, fn in (("NT", lambda: ops.gemm(a, w, trans_b=True, out=c)),
                             ("NN", lambda: ops.gemm(a, wt, out=c)),
                             ("NT+acc", lambda: ops.gemm(a, w, trans_b=True, out=c, accumulate=True))):
                ms = timeit(fn)
                print(f"{name:8s} {M:7d} {N:5d} {K:5d} {ms:8.3f} {flops / ms / 1e9:7.1f} {byts / ms / 1e6:7.0f}")
            dy = torch.randn(M, N, device=dev)
            ms = timeit(lambda: ops.gemm(dy, a, trans_a=True))
            print(f"{'TN(dW)':8s} {M:7d} {N:5d} {K:5d} {ms:8.3f} {flops / ms / 1e9:7.1f} {4.0 * (M * K + M * N) / ms / 1e6:7.0f}")
            del a, w, wt, c, dy


if __name__ == "__main__":
    main()
