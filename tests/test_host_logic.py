"""Host-side logic that needs no GPU: the per-batch code ordering behind gnnx_code_segment_sum_sorted (pure torch ops, so
it runs on CPU tensors too), the launch counter's bookkeeping of graph replays, and the environment switches' defaults."""
import numpy as np
import pytest
import torch

from gnnepcsaft_b200 import ops, streams


@pytest.mark.parametrize("n,ncodes,seed", [(0, 60, 0), (1, 60, 1), (255, 60, 2), (256, 60, 3), (257, 60, 4), (10000, 60, 5), (5000, 7, 6)])
def test_code_order_is_a_stable_sort_with_consistent_pointers(n, ncodes, seed):
    rng = np.random.default_rng(seed)
    codes = torch.from_numpy(rng.integers(0, ncodes, n).astype(np.uint8))
    if n > 10:
        codes[: n // 2] = 3                                  # one dominant code, several empty ones
    ops._code_order_cache.clear()
    order, code_ptr, chunk_ptr = ops.code_order(codes, ncodes)
    assert order.dtype == torch.int32 and code_ptr.dtype == torch.int32 and chunk_ptr.dtype == torch.int32
    assert order.shape == (n,) and code_ptr.shape == (ncodes + 1,) and chunk_ptr.shape == (ncodes + 1,)
    want = np.argsort(codes.numpy(), kind="stable")
    assert np.array_equal(order.numpy(), want)
    cnt = np.bincount(codes.numpy(), minlength=ncodes)
    assert np.array_equal(code_ptr.numpy(), np.concatenate([[0], np.cumsum(cnt)]))
    assert np.array_equal(chunk_ptr.numpy(), np.concatenate([[0], np.cumsum((cnt + 255) // 256)]))
    # the kernel's launch bound: sum_k ceil(n_k / 256) <= n / 256 + ncodes
    assert int(chunk_ptr[-1]) <= n // 256 + ncodes
    # cached on the tensor: the same object comes back, a changed tensor does not reuse it
    assert ops.code_order(codes, ncodes)[0] is order
    if n:
        codes[0] = (int(codes[0]) + 1) % ncodes              # in-place edit bumps the version
        assert ops.code_order(codes, ncodes)[0] is not order


def test_switch_defaults():
    # what the measured numbers were taken with (INTEGRATION.md lists the switches)
    import os
    if "GNNX_PREP_AHEAD" not in os.environ and "GNNX_PREP_GRAPH" not in os.environ:
        assert streams.PREP_AHEAD and streams.PREP_GRAPH
    assert ops.CODE_SORTED_MIN_ROWS == 8192
