"""gnnx_gemm_chain: Linear -> ReLU -> Linear (-> ...) chains in one launch (k_gemm_tf32_chain).  Every layer's output and
packed ReLU mask must equal what ``ops.gemm`` produces layer by layer on the same tier -- bit for bit: same tiles, same
K order, same epilogue -- for row counts that leave partial groups / partial tiles / idle clusters, and the fp32 tier
(served layer by layer) must agree with an fp64 product."""
import pytest
import torch

from gnnepcsaft_b200 import ops
from gnnepcsaft_b200._lib import lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def _tier():
    ops.set_matmul_precision("tf32")
    ops.FORCE_TC = True
    torch.manual_seed(0)
    yield
    ops.FORCE_TC = False
    ops.set_matmul_precision("fp32")


@pytest.mark.parametrize("M,W,n", [(768, 64, 2), (1000, 128, 3), (5000, 256, 3), (40000, 256, 3), (101887, 256, 3),
                                    (3333, 96, 4), (205206, 256, 2), (130 * 256 + 17, 32, 3)])
def test_chain_equals_layer_by_layer(M, W, n):
    x = torch.randn(M, W, device=DEV)
    layers = [((torch.randn(W, W, device=DEV) / W ** 0.5), torch.randn(W, device=DEV) * 0.1, p < n - 1, p < n - 1) for p in range(n)]
    tc0 = lib().tc_launch_count()
    outs, bits = ops.gemm_chain(x, layers)
    assert lib().tc_launch_count() - tc0 == 1, "one tcgen05 launch for the whole chain"
    a = x
    for p, (w, b, relu, want) in enumerate(layers):
        if want:
            ref, rbits = ops.gemm(a, w, trans_b=True, bias=b, relu=relu, want_relu_bits=True)
            assert torch.equal(bits[p], rbits), f"mask of layer {p}"
        else:
            ref = ops.gemm(a, w, trans_b=True, bias=b, relu=relu)
            assert bits[p] is None
        assert torch.equal(outs[p], ref), f"layer {p}: max|d| {float((outs[p] - ref).abs().max()):.3e}"
        a = ref
    again, _ = ops.gemm_chain(x, layers)
    assert all(torch.equal(u, v) for u, v in zip(outs, again))            # deterministic


def test_chain_on_the_fp32_tier_and_small_inputs_is_served_layer_by_layer():
    ops.set_matmul_precision("fp32")
    ops.FORCE_TC = False
    x = torch.randn(300, 64, device=DEV)
    layers = [(torch.randn(64, 64, device=DEV) / 8, torch.randn(64, device=DEV), True, True),
              (torch.randn(64, 64, device=DEV) / 8, None, False, False)]
    tc0 = lib().tc_launch_count()
    outs, bits = ops.gemm_chain(x, layers)
    assert lib().tc_launch_count() == tc0
    h = torch.relu(x.double() @ layers[0][0].double().t() + layers[0][1].double())
    assert torch.allclose(outs[0].double(), h, rtol=1e-5, atol=1e-5)
    assert torch.allclose(outs[1].double(), outs[0].double() @ layers[1][0].double().t(), rtol=1e-5, atol=1e-5)
    assert bits[0] is not None and bits[1] is None
