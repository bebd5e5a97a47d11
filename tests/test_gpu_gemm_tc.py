"""tcgen05 (kind::tf32) GEMM kernels against an fp64 product of the same operands.

Tolerance (stated separately from the fp32 tier, north_star): operands are rounded to tf32 (10-bit mantissa,
unit roundoff 2^-11) and accumulated in fp32, so |C - C64| <= ~2 * 2^-11 * (|A| |B|) elementwise; the test
bounds the error by 3 * 2^-11 * (|A| @ |B|).  The layouts under test are the three the hot path uses:
A K-major x B K-major (Linear forward), A K-major x B MN-major (dX = dY W) and both MN-major (dW = dY^T X).
"""
import pytest
import torch

from gnnepcsaft_b200 import ops
from gnnepcsaft_b200._lib import lib

pytestmark = pytest.mark.gpu
DEV = "cuda"
U_TF32 = 2.0 ** -11


def _check(out, a64, b64, extra64=None, relu=False, what=""):
    ref = a64 @ b64
    if extra64 is not None:
        ref = ref + extra64
    bound = 3 * U_TF32 * (a64.abs() @ b64.abs()) + 1e-6
    if relu:
        ref = ref.clamp(min=0)
    err = (out.double().cpu() - ref).abs()
    assert bool(torch.isfinite(out).all()), f"{what}: non-finite"
    worst = float((err / bound).max())
    assert worst <= 1.0, f"{what}: error/bound = {worst:.3g} (max abs err {float(err.max()):.3g})"
    # and it must really be a tf32-class result, not garbage that happens to be small: relative Frobenius error
    rel = float(err.norm() / ref.norm().clamp(min=1e-30))
    assert rel < 2e-3, f"{what}: relative Frobenius error {rel:.3g}"


@pytest.fixture(autouse=True)
def _prec():
    ops.set_matmul_precision("tf32")
    torch.manual_seed(0)
    yield
    ops.set_matmul_precision("fp32")


def _tc_taken(fn):
    """Run fn with the size heuristic off (ops.FORCE_TC: a product the tcgen05 kernels cannot take raises instead of
    falling back) and assert that a tensor-core kernel really was launched (gnnx_tc_launch_count)."""
    n0 = lib().tc_launch_count()
    ops.FORCE_TC = True
    try:
        out = fn()
    finally:
        ops.FORCE_TC = False
    assert lib().tc_launch_count() > n0, "no tcgen05 kernel was launched"
    return out


@pytest.mark.parametrize("M,N,K", [(256, 256, 256), (300, 256, 256), (1000, 768, 1024), (4096, 128, 256), (257, 64, 96),
                                   (5000, 48, 40), (100000, 256, 256), (777, 320, 72)])
def test_linear_forward_layout(M, N, K):
    a, w, bias = torch.randn(M, K), torch.randn(N, K), torch.randn(N)
    out = _tc_taken(lambda: ops.gemm(a.to(DEV), w.to(DEV), trans_b=True, bias=bias.to(DEV)))
    _check(out, a.double(), w.double().t(), bias.double(), what=f"NT {M}x{N}x{K}")
    out = _tc_taken(lambda: ops.gemm(a.to(DEV), w.to(DEV), trans_b=True, bias=bias.to(DEV), relu=True))
    _check(out, a.double(), w.double().t(), bias.double(), relu=True, what=f"NT+relu {M}x{N}x{K}")


def test_forward_strided_views_and_accumulate():
    M = 3000
    x = torch.randn(M, 1024)
    w = torch.randn(256, 1300)
    out = torch.randn(M, 768)
    ref_prev = out[:, 256:512].double().clone()
    xd, wd, od = x.to(DEV), w.to(DEV), out.to(DEV)
    ops.gemm(xd[:, 256:768], wd[:, 260:772], trans_b=True, out=od[:, 256:512], accumulate=True)
    _check(od[:, 256:512], x[:, 256:768].double(), w[:, 260:772].double().t(), ref_prev, what="strided accumulate")
    assert torch.equal(od[:, :256].cpu(), out[:, :256]) and torch.equal(od[:, 512:].cpu(), out[:, 512:])


@pytest.mark.parametrize("M,N,K", [(1000, 256, 256), (3000, 1024, 768), (4096, 64, 128), (100000, 256, 256), (513, 96, 200)])
def test_dx_layout(M, N, K):
    dy, w = torch.randn(M, K), torch.randn(K, N)          # dX = dY @ W, W is [out=K, in=N] row-major
    out = _tc_taken(lambda: ops.gemm(dy.to(DEV), w.to(DEV)))
    _check(out, dy.double(), w.double(), what=f"NN {M}x{N}x{K}")


@pytest.mark.parametrize("Mo,No,R", [(256, 256, 5000), (768, 1024, 3001), (128, 256, 4096), (256, 128, 100000), (64, 96, 777),
                                     (384, 288, 2000)])
def test_dw_layout(Mo, No, R):
    dy, x = torch.randn(R, Mo), torch.randn(R, No)
    out = _tc_taken(lambda: ops.gemm(dy.to(DEV), x.to(DEV), trans_a=True))
    _check(out, dy.double().t(), x.double(), what=f"TN {Mo}x{No}x{R}")
    again = _tc_taken(lambda: ops.gemm(dy.to(DEV), x.to(DEV), trans_a=True))
    assert torch.equal(out, again), "dW must be deterministic (ordered split reduction, no atomics)"


def test_dw_strided_accumulate():
    R = 20000
    dy, x = torch.randn(R, 768), torch.randn(R, 1024)
    dw = torch.randn(256, 1300)
    prev = dw[:, 100:356].double().clone()
    dwd = dw.to(DEV)
    ops.gemm(dy.to(DEV)[:, 256:512], x.to(DEV)[:, 512:768], trans_a=True, out=dwd[:, 100:356], accumulate=True)
    _check(dwd[:, 100:356], dy[:, 256:512].double().t(), x[:, 512:768].double(), prev, what="dW strided accumulate")
    assert torch.equal(dwd[:, :100].cpu(), dw[:, :100]) and torch.equal(dwd[:, 356:].cpu(), dw[:, 356:])


def test_small_or_unaligned_shapes_fall_back_to_fp32_kernel():
    a, w = torch.randn(60, 13), torch.randn(64, 13)        # the 60-row bond table product: SIMT path, fp32-exact class
    t0 = lib().tc_launch_count()
    out = ops.gemm(a.to(DEV), w.to(DEV), trans_b=True)
    assert lib().tc_launch_count() == t0
    ref = a.double() @ w.double().t()
    assert float((out.double().cpu() - ref).abs().max()) < 1e-4
    ops.FORCE_TC = True                                    # and pinning it onto the tensor path must fail loudly
    try:
        with pytest.raises(RuntimeError):
            ops.gemm(a.to(DEV), w.to(DEV), trans_b=True)
    finally:
        ops.FORCE_TC = False


@pytest.mark.parametrize("prec", ["tf32", "fp32"])
def test_fused_relu_mask_and_accumulate(prec):
    """dX = (dY W) * (y_in > 0) with the mask applied in the epilogue, alone and on top of accumulate."""
    ops.set_matmul_precision(prec)
    M, K, N = 3000, 256, 512
    dy, w, y = torch.randn(M, K), torch.randn(K, N), torch.randn(M, N)
    ref = (dy.double() @ w.double()) * (y.double() > 0)
    out = ops.gemm(dy.to(DEV), w.to(DEV), relu_mask=y.to(DEV))
    err = (out.double().cpu() - ref).abs()
    bound = 3 * U_TF32 * (dy.double().abs() @ w.double().abs()) + 1e-6
    assert float((err / bound).max()) <= 1.0
    assert bool((out.cpu()[y <= 0] == 0).all())
    prev = torch.randn(M, N)
    out2 = prev.to(DEV).clone()
    ops.gemm(dy.to(DEV), w.to(DEV), out=out2, accumulate=True, relu_mask=y.to(DEV))
    ref2 = (dy.double() @ w.double() + prev.double()) * (y.double() > 0)
    assert float(((out2.double().cpu() - ref2).abs() / bound).max()) <= 1.0


def _pack_bits(y: torch.Tensor) -> torch.Tensor:
    """Reference packing: bit c%32 of word [r, c//32] = (y[r, c] > 0); int32 words as the kernels store them."""
    M, N = y.shape
    words = (N + 31) // 32
    pad = torch.zeros(M, words * 32, dtype=torch.bool)
    pad[:, :N] = y > 0
    w = (pad.view(M, words, 32).to(torch.int64) << torch.arange(32, dtype=torch.int64)).sum(-1)
    return torch.where(w >= 2 ** 31, w - 2 ** 32, w).to(torch.int32)


@pytest.mark.parametrize("prec", ["tf32", "fp32"])
@pytest.mark.parametrize("M,K,N", [(3000, 256, 256), (1000, 64, 96), (333, 128, 512), (5000, 256, 48)])
def test_packed_relu_bits_produced_and_consumed(prec, M, K, N):
    """A ReLU'd forward product emits the 1-bit mask of its own output; the matching dX product consumes it
    (GNNX_GEMM_MASK_BITS) and must equal the fp32-mask variant bit for bit."""
    ops.set_matmul_precision(prec)
    torch.manual_seed(M + N)
    x, w, b = torch.randn(M, K), torch.randn(N, K) / K ** 0.5, torch.randn(N)
    y, bits = ops.gemm(x.to(DEV), w.to(DEV), trans_b=True, bias=b.to(DEV), relu=True, want_relu_bits=True)
    words = (N + 31) // 32
    assert bits.shape == (M, words) and bits.dtype == torch.int32
    want = _pack_bits(y.cpu())
    valid = N - 32 * (words - 1)                       # bits past column N in the last word are don't-care
    keep = torch.full((words,), -1, dtype=torch.int64)
    keep[-1] = (1 << valid) - 1
    assert torch.equal(bits.cpu().to(torch.int64) & keep, want.to(torch.int64) & keep), "packed mask differs from (y > 0)"
    dy, w2 = torch.randn(M, K), torch.randn(K, N)
    a = ops.gemm(dy.to(DEV), w2.to(DEV), relu_mask=y)
    bq = ops.gemm(dy.to(DEV), w2.to(DEV), relu_mask=bits)
    assert torch.equal(a, bq), "packed-mask dX differs from the fp32-mask dX"
    prev = torch.randn(M, N)
    o1, o2 = prev.to(DEV).clone(), prev.to(DEV).clone()
    ops.gemm(dy.to(DEV), w2.to(DEV), out=o1, accumulate=True, relu_mask=y)
    ops.gemm(dy.to(DEV), w2.to(DEV), out=o2, accumulate=True, relu_mask=bits)
    assert torch.equal(o1, o2), "packed-mask accumulate variant differs"


@pytest.mark.parametrize("prec", ["tf32", "fp32"])
@pytest.mark.parametrize("M,K,N,tb", [(101028, 256, 256, False), (3000, 256, 512, False), (1000, 64, 96, True),
                                      (333, 128, 256, False), (130, 256, 256, False), (5000, 768, 256, False),
                                      (60, 256, 64, False), (4096, 32, 64, True)])
def test_fused_column_sums_of_a(prec, M, K, N, tb):
    """``want_colsum_a``: the bias gradient colsum(dY) delivered by the dX product itself (summed from the staged A
    tiles on the tensor-core path for K <= 256, by the two-pass kernel otherwise); the product is unchanged."""
    ops.set_matmul_precision(prec)
    torch.manual_seed(M + K)
    a = torch.randn(M, K) + 0.25
    b = torch.randn(N, K) if tb else torch.randn(K, N)
    plain = ops.gemm(a.to(DEV), b.to(DEV), trans_b=tb)
    out, cs = ops.gemm(a.to(DEV), b.to(DEV), trans_b=tb, want_colsum_a=True)
    assert torch.equal(out, plain)
    ref = a.double().sum(0)
    scale = a.double().abs().sum(0)
    u = U_TF32 if prec == "tf32" else 2.0 ** -24                 # tf32 tier: operands as the tensor pipe sees them
    assert float(((cs.double().cpu() - ref).abs() / (scale * u * 2 + 1e-6)).max()) <= 1.0
    y = torch.randn(M, N)
    prev = torch.randn(M, N)
    o2 = prev.to(DEV).clone()
    _, cs2 = ops.gemm(a.to(DEV), b.to(DEV), trans_b=tb, out=o2, accumulate=True, relu_mask=y.to(DEV), want_colsum_a=True)
    assert torch.equal(cs2, cs), "column sums must not depend on the epilogue variant"
    o3 = prev.to(DEV).clone()
    ops.gemm(a.to(DEV), b.to(DEV), trans_b=tb, out=o3, accumulate=True, relu_mask=y.to(DEV))
    assert torch.equal(o2, o3)


@pytest.mark.parametrize("n,R,Mo,No", [(3, 5000, 256, 256), (2, 70001, 128, 256), (4, 3000, 256, 128), (3, 1000, 96, 200),
                                       (1, 4096, 256, 256)])
def test_batched_weight_gradients(n, R, Mo, No):
    """gnnx_gemm_dw_batched: n products dY_z^T X_z of one shape in one launch (blockIdx.z = product) must each meet the
    tf32 bound, be bit-identical across reruns and independent of their batch mates; shapes the batched kernel does not
    take (96 x 200: not multiples of 32; n = 1) are served one by one with exactly gemm(trans_a=True)'s result."""
    dys = [torch.randn(R, Mo, device=DEV) for _ in range(n)]
    xs = [torch.randn(R, No, device=DEV) for _ in range(n)]
    outs = ops.gemm_dw_batched(dys, xs)
    again = ops.gemm_dw_batched(dys, xs)
    for z in range(n):
        _check(outs[z], dys[z].double().cpu().t(), xs[z].double().cpu(), what=f"batched dW {z}")
        assert torch.equal(outs[z], again[z])
    if n >= 2:
        other = ops.gemm_dw_batched(dys[::-1], xs[::-1])[::-1]         # same products at other batch positions
        for z in range(n):
            assert torch.equal(outs[z], other[z])
    if Mo % 32 or No % 32 or n == 1:
        for z in range(n):
            assert torch.equal(outs[z], ops.gemm(dys[z], xs[z], trans_a=True))
    fp32 = ops.gemm_dw_batched(dys, xs, precision=ops.PREC_FP32)
    for z in range(n):
        ref = dys[z].double().t() @ xs[z].double()
        assert float((fp32[z].double() - ref).abs().max()) <= 1e-5 * float(ref.abs().max()) + 1e-4


def test_batched_weight_gradients_strided_operands():
    """Operands that are column slices of wider matrices (row stride != width), as the packed layouts hand them over."""
    R = 9000
    big_dy, big_x = torch.randn(R, 640, device=DEV), torch.randn(R, 512, device=DEV)
    dys = [big_dy[:, 0:256], big_dy[:, 256:512]]
    xs = [big_x[:, 256:512], big_x[:, 0:256]]
    outs = ops.gemm_dw_batched(dys, xs)
    for z in range(2):
        _check(outs[z], dys[z].double().cpu().t(), xs[z].double().cpu(), what=f"strided batched dW {z}")
