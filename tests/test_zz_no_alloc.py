"""gnnx_colsum_noalloc: the column sums of gnnx_colsum with the chunk partials in the library's own file-scope device
array (no caller workspace, no cudaMalloc / cudaFree / torch allocation inside the call).  Called through the C ABI with
raw pointers and compared with the existing entry point on two calls in a row -- the second on a smaller, unrelated
input, so partials the first call left in the shared array would show up.  Collected last on purpose (file name).

No counterpart in the reference (include/gnnx.h).
"""
import ctypes

import pytest
import torch

from gnnepcsaft_b200._lib import GnnxError, lib
from tests.util import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda"
RTOL, ATOL = 1e-5, 1e-6            # the fp32 tier's tolerance (tests/test_gpu_ops.py); the two entry points are bit-identical


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _colsum_existing(x, out):
    L = lib()
    rows, cols = x.shape
    nb = L.query("gnnx_colsum_workspace_bytes", rows, cols)
    ws = torch.empty(max(nb, 1), dtype=torch.uint8, device=x.device)
    L.call("gnnx_colsum", x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), ws.data_ptr(), nb, _stream())


def _colsum_noalloc(x, out):
    rows, cols = x.shape
    lib().call("gnnx_colsum_noalloc", x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), _stream())


def test_colsum_noalloc_matches_colsum_on_two_calls_in_a_row():
    torch.manual_seed(7)
    L = lib()
    assert L.query("gnnx_colsum_noalloc_max_bytes") == 1 << 20
    # first call: cfg2-sized [N, H] (float4 kernel, the full chunk count); second: fewer rows and columns, a strided
    # view with an odd column count (scalar kernel) -- the shared array still holds the first call's partials
    big = torch.randn(101_887, 256, device=DEV)
    small = torch.randn(3_001, 203, device=DEV)[:, 5:74]
    inputs = [big, small]
    want = [torch.empty(x.shape[1], device=DEV) for x in inputs]
    got = [torch.full((x.shape[1],), float("nan"), device=DEV) for x in inputs]
    for x, w in zip(inputs, want):
        _colsum_existing(x, w)
    torch.cuda.synchronize()
    mallocs = torch.cuda.memory_stats()["allocation.all.allocated"]
    free0, _ = torch.cuda.mem_get_info()
    for x, g in zip(inputs, got):                    # two calls in a row, nothing in between
        _colsum_noalloc(x, g)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == mallocs       # no torch allocation
    assert free1 == free0                                                         # no cudaMalloc behind torch's back
    for x, g, w in zip(inputs, got, want):
        assert_close(g, w, RTOL, ATOL, f"colsum_noalloc {tuple(x.shape)} vs colsum")
        assert torch.equal(g, w), "the two entry points run the same kernels in the same order"
        assert_close(g, x.double().sum(0), RTOL, ATOL, f"colsum_noalloc {tuple(x.shape)} vs fp64")
    # and once more on the first input: same bits again
    again = torch.empty(256, device=DEV)
    _colsum_noalloc(big, again)
    assert torch.equal(again, want[0])


def test_colsum_noalloc_edge_shapes_and_limit():
    L = lib()
    out = torch.full((16,), 3.0, device=DEV)
    _colsum_noalloc(torch.empty(0, 16, device=DEV), out)                         # zero rows: zeros
    assert float(out.abs().sum()) == 0.0
    one = torch.randn(1, 16, device=DEV)
    _colsum_noalloc(one, out)
    assert torch.equal(out, one[0])
    # beyond the built-in scratch: [1, 300 000] needs 1.2 MB of partials > 1 MiB -> a clear error, output untouched
    wide = torch.zeros(1, 300_000, device=DEV)
    o = torch.full((300_000,), 5.0, device=DEV)
    assert L.query("gnnx_colsum_workspace_bytes", 1, 300_000) > L.query("gnnx_colsum_noalloc_max_bytes")
    with pytest.raises(GnnxError, match="built-in scratch"):
        _colsum_noalloc(wide, o)
    torch.cuda.synchronize()
    assert float(o.min()) == 5.0
