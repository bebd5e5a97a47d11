"""Per-shape timing of the tensor-core row-tile products of a cfg2 step, in isolation (inputs rotate over buffers larger
than L2, CUDA events around 20 back-to-back launches), with the CTA-pair kernel off and on.  Prints us per launch and
the fraction of the measured HBM peak on the algorithmic bytes."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gnnepcsaft_b200 import ops  # noqa: E402

DEV = "cuda"
N_ROWS, E_ROWS, CAP = 101887, 205206, 104448
SHAPES = [  # name, M, N, K, trans_b, masked + colsum
    ("fwd  N x 256 x 256", N_ROWS, 256, 256, True, False),
    ("fwd  pq N x 512 x 256", N_ROWS, 512, 256, True, False),
    ("fwd  pre E x 256 x 256", E_ROWS, 256, 256, True, False),
    ("fwd  grouped-like cap x 256 x 1280", CAP, 256, 1280, True, False),
    ("dX   N x 256 x 256", N_ROWS, 256, 256, False, True),
    ("dX   E x 256 x 256", E_ROWS, 256, 256, False, True),
    ("dX   dagg-like cap x 1280 x 256", CAP, 1280, 256, False, True),
    ("dX   dpq N x 256 x 512", N_ROWS, 256, 512, False, False),
    ("dX   N x 256 x 256, mask only", N_ROWS, 256, 256, False, "mask"),
    ("dX   N x 256 x 256, colsum only", N_ROWS, 256, 256, False, "colsum"),
    ("dX   N x 256 x 256, plain", N_ROWS, 256, 256, False, False),
    ("fwd  N x 256 x 256, bias relu bits", N_ROWS, 256, 256, True, "fwdbits"),
]


def main():
    ops.set_matmul_precision("tf32")
    peak = 6552.6
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    reps, nbuf = 20, 3
    for name, M, N, K, tb, masked in SHAPES:
        a = [torch.randn(M, K, device=DEV) for _ in range(nbuf)]
        c = [torch.empty(M, N, device=DEV) for _ in range(nbuf)]
        w = torch.randn(N, K, device=DEV) if tb else torch.randn(K, N, device=DEV)
        bits = (torch.rand(M, (N + 31) // 32, device=DEV) * 2 ** 31).to(torch.int32) if masked else None
        bias = torch.randn(N, device=DEV)
        gb = 4.0 * M * (N + K) / 1e9
        row = f"{name:40s} {gb * 1e3:7.1f} MB "
        for pair in os.environ.get("SWEEP", "0,1").split(","):
            os.environ[os.environ.get("SWEEP_VAR", "GNNX_GEMM_PAIR")] = pair
            def run(i):
                if masked == "mask":
                    ops.gemm(a[i % nbuf], w, trans_b=tb, relu_mask=bits, out=c[i % nbuf])
                elif masked == "colsum":
                    ops.gemm(a[i % nbuf], w, trans_b=tb, want_colsum_a=True, out=c[i % nbuf])
                elif masked == "fwdbits":
                    ops.gemm(a[i % nbuf], w, trans_b=tb, bias=bias, relu=True, want_relu_bits=True, out=c[i % nbuf])
                elif masked and K <= 256:
                    ops.gemm(a[i % nbuf], w, trans_b=tb, relu_mask=bits, want_colsum_a=True, out=c[i % nbuf])
                else:
                    ops.gemm(a[i % nbuf], w, trans_b=tb, out=c[i % nbuf])
            for i in range(3):
                run(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for i in range(reps):
                run(i)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / reps
            row += f"| pair={pair}: {us:7.1f} us  {gb / (us * 1e-6) / peak:5.2f} of HBM "
        print(row)


if __name__ == "__main__":
    main()
