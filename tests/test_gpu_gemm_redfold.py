"""Weight-gradient products C = A^T B on the tensor cores: the ordered fold of the row-split partials inside the product's
own launch (grid barrier, every CTA sums a slice in split order) against the separate k_splitk_reduce launch
(GNNX_RED_FOLD=0).  Same partials, same order of the sum: the results must be BIT-IDENTICAL -- single products, with
bias / accumulate, and the batched form of a PNAConv layer's chain."""
import os

import pytest
import torch

from gnnepcsaft_b200 import ops
from gnnepcsaft_b200._lib import lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def _tier():
    was = os.environ.get("GNNX_RED_FOLD")
    ops.set_matmul_precision("tf32")
    ops.FORCE_TC = True
    torch.manual_seed(0)
    yield
    ops.FORCE_TC = False
    ops.set_matmul_precision("fp32")
    if was is None:
        os.environ.pop("GNNX_RED_FOLD", None)
    else:
        os.environ["GNNX_RED_FOLD"] = was


def _both(fn):
    out = []
    for fold in ("0", "1"):
        os.environ["GNNX_RED_FOLD"] = fold
        tc0 = lib().tc_launch_count()
        out.append(fn())
        assert lib().tc_launch_count() > tc0, "the tensor-core kernel must have run"
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("R,Mo,No", [(101887, 256, 256), (205206, 256, 256), (40000, 512, 256), (5000, 64, 32), (300, 128, 96),
                                      (70000, 256, 1280), (33333, 32, 256)])
def test_fold_in_kernel_equals_separate_reduce(R, Mo, No):
    a = torch.randn(R, Mo, device=DEV)
    b = torch.randn(R, No, device=DEV)
    plain = _both(lambda: ops.gemm(a, b, trans_a=True))
    assert torch.equal(*plain), f"max|d| {float((plain[0] - plain[1]).abs().max()):.3e}"
    ref = a.double().t() @ b.double()
    bound = 3 * 2.0 ** -11 * float((a.abs().double().t() @ b.abs().double()).max())
    assert float((plain[1].double() - ref).abs().max()) <= bound
    base = torch.randn(Mo, No, device=DEV)
    acc = _both(lambda: ops.gemm(a, b, trans_a=True, out=base.clone(), accumulate=True))
    assert torch.equal(*acc)
    again = _both(lambda: ops.gemm(a, b, trans_a=True))
    assert torch.equal(plain[1], again[1]), "rerun (the barrier counters must be back at zero)"


@pytest.mark.parametrize("n", [2, 3, 4])
def test_batched_weight_gradients(n):
    R, W = 50000, 256
    dys = [torch.randn(R, W, device=DEV) for _ in range(n)]
    xs = [torch.randn(R, W, device=DEV) for _ in range(n)]
    got = _both(lambda: ops.gemm_dw_batched(dys, xs))
    for z in range(n):
        assert torch.equal(got[0][z], got[1][z]), f"problem {z}"
