"""k_gemm_tf32_pair (GNNX_GEMM_PAIR=1): the row-tile products on CTA pairs -- one tcgen05.mma.cta_group::2 per k-step for
two SMs, half of the weight tile in each CTA's shared memory, six 32 KB stages instead of four 48 KB ones.  Same tiles,
same K order and same epilogue as k_gemm_tf32_rows2, so every output (product, packed ReLU mask, fused column sums,
grouped products) must be BIT-IDENTICAL with the pair kernel on and off; the product itself is pinned against fp64 by
tests/test_gpu_gemm_tc.py on whichever kernel is the default."""
import os

import numpy as np
import pytest
import torch

from gnnepcsaft_b200 import graph, ops
from gnnepcsaft_b200._lib import lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def _tier():
    was = os.environ.get("GNNX_GEMM_PAIR")
    ops.set_matmul_precision("tf32")
    ops.FORCE_TC = True
    torch.manual_seed(0)
    yield
    ops.FORCE_TC = False
    ops.set_matmul_precision("fp32")
    if was is None:
        os.environ.pop("GNNX_GEMM_PAIR", None)
    else:
        os.environ["GNNX_GEMM_PAIR"] = was


def _both(fn):
    out = []
    for pair in ("0", "1"):
        os.environ["GNNX_GEMM_PAIR"] = pair
        tc0 = lib().tc_launch_count()
        out.append(fn())
        assert lib().tc_launch_count() > tc0
    torch.cuda.synchronize()
    return out


def _same(a, b, what):
    if a is None or b is None:
        assert a is None and b is None, what
    elif isinstance(a, (tuple, list)):
        for i, (u, v) in enumerate(zip(a, b)):
            _same(u, v, f"{what}[{i}]")
    else:
        assert torch.equal(a, b), f"{what}: max|d| {float((a.float() - b.float()).abs().max()):.3e}"


@pytest.mark.parametrize("M,N,K", [(256, 256, 256), (257, 64, 32), (1000, 128, 96), (5000, 256, 256), (101887, 256, 256),
                                    (40000, 512, 256), (33000, 256, 1280), (130 * 256 + 1, 192, 200), (70000, 1280, 256)])
@pytest.mark.parametrize("trans_b", [True, False])
def test_pair_equals_rows2(M, N, K, trans_b):
    if not trans_b and N % 64:
        pytest.skip("MN-major weights come in 32-column chunks; pairs need two of them")
    a = torch.randn(M, K, device=DEV)
    w = torch.randn(N, K, device=DEV) / K ** 0.5 if trans_b else torch.randn(K, N, device=DEV) / K ** 0.5
    b = torch.randn(N, device=DEV)
    plain = _both(lambda: ops.gemm(a, w, trans_b=trans_b))
    _same(*plain, "product")
    ref = a.double() @ (w.double().t() if trans_b else w.double())
    assert float((plain[1].double() - ref).abs().max()) <= 3 * 2.0 ** -11 * float((a.abs().double() @ (w.abs().double().t() if trans_b else w.abs().double())).max())
    _same(*_both(lambda: ops.gemm(a, w, trans_b=trans_b, bias=b, relu=True, want_relu_bits=True)), "bias + relu + packed mask")
    if K <= 256:
        bits = (torch.rand(M, (N + 31) // 32, device=DEV) * 2 ** 31).to(torch.int32)
        _same(*_both(lambda: ops.gemm(a, w, trans_b=trans_b, relu_mask=bits, want_colsum_a=True)), "masked product + column sums of A")
    again = _both(lambda: ops.gemm(a, w, trans_b=trans_b))
    _same(plain[1], again[1], "rerun")


@pytest.mark.parametrize("N,maxdeg,H,Ho", [(3000, 4, 128, 128), (20000, 4, 256, 256), (1200, 9, 64, 64)])
def test_pair_grouped_products(N, maxdeg, H, Ho):
    rng = np.random.default_rng(N)
    deg = rng.integers(0, maxdeg + 1, N)
    deg[0] = maxdeg
    rowptr = torch.from_numpy(np.concatenate([[0], np.cumsum(deg)]).astype(np.int32)).to(DEV)
    bk = graph.DegreeBuckets(rowptr, maxdeg + 1)
    D = bk.n_groups
    a = torch.randn(bk.capacity, 5 * H, device=DEV)
    weff = torch.randn(D * Ho, 5 * H, device=DEV) / (5 * H) ** 0.5
    bias = torch.randn(Ho, device=DEV)
    _same(*_both(lambda: ops.gemm_grouped(a, weff, bk, trans_b=True, bias=bias, relu=True, want_relu_bits=True)), "grouped forward")
    dy = torch.randn(bk.capacity, Ho, device=DEV)
    _same(*_both(lambda: ops.gemm_grouped(dy, weff, bk, trans_b=False, want_colsum_a=True)), "grouped dX + column sums")
