"""Tensor-level wrappers over the C ABI (``include/gnnx.h``).

Every function validates device / dtype / contiguity in Python, allocates outputs and workspaces
with torch (plumbing only), and launches on torch's current CUDA stream.  No function here has a
CPU path: a CPU tensor raises.
"""

from __future__ import annotations

import os

from typing import Optional, Tuple

import torch
from torch import Tensor

from ._lib import lib

PREC_FP32, PREC_TF32 = 0, 1
GEMM_RELU, GEMM_ACCUMULATE, GEMM_MASK_BITS, GEMM_FORCE_TC = 1, 2, 4, 8
# tests set this to pin every tf32-tier product onto the tcgen05 kernels (products they cannot take then raise)
FORCE_TC = False
# PNAConv forward: aggregate -> first post Linear as ONE tcgen05 kernel whose A operand is produced in shared memory
# (csrc/gemm_tc.cu: k_pna_post_fused).  Correct (tests/test_gpu_pna_fused.py) but OFF by default: on a B200 it takes 674 us
# per cfg2 layer against 156 + 139 us for the aggregate kernel plus the grouped product it replaces (DESIGN.md section 8
# has the ablation: a software agent in the stage hand-shake alone costs 150 us).  GNNX_FUSED_POST=1 switches it on.
FUSED_POST = os.environ.get("GNNX_FUSED_POST", "0") == "1"
# torch's names map onto the two tiers that exist: "highest" = CUDA-core fp32, "high" = tcgen05 tf32.  The reference's
# "medium" (bf16-class products, train/train.py:54) has no kernels of its own here: it is served by the tf32 tier, which
# is MORE precise than requested, and results are reported as tf32 (bench.py dtype), never as bf16.
_PRECISIONS = {"fp32": PREC_FP32, "highest": PREC_FP32, "tf32": PREC_TF32, "high": PREC_TF32, "medium": PREC_TF32}

_default_precision = PREC_FP32


def set_matmul_precision(name: str) -> None:
    """Mirror of ``torch.set_float32_matmul_precision`` for the gnnx Linear layers
    (reference: train/train.py:54 uses "medium"; parity tests use "highest")."""
    global _default_precision
    if name not in _PRECISIONS:
        raise ValueError(f"gnnx: unknown matmul precision {name!r} (available: {sorted(_PRECISIONS)}; no bf16 tier is built)")
    _default_precision = _PRECISIONS[name]


def get_matmul_precision() -> int:
    return _default_precision


_raw_stream = torch._C._cuda_getCurrentRawStream


def _stream() -> int:
    # torch.cuda.current_stream() builds a Python Stream object (~15 us); the raw handle is what the C ABI wants
    return _raw_stream(torch.cuda.current_device())


def _chk(t: Tensor, dtype, name: str) -> Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"gnnx: `{name}` must be a CUDA tensor (no CPU fallback exists)")
    if t.dtype != dtype:
        raise TypeError(f"gnnx: `{name}` must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"gnnx: `{name}` must be contiguous")
    return t


def _p(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


_WS_CACHE: dict = {}
_WS_QUERY_CACHE: dict = {}


def _ws(nbytes: int, device) -> Tuple[Optional[Tensor], int]:
    """Scratch for one call.  Kernels of one stream run in order, so a single grow-only buffer per (device, stream)
    is reused by every call instead of a fresh torch.empty each time."""
    if nbytes == 0:
        return None, 0
    key = (device.index if device.index is not None else torch.cuda.current_device(), _stream())
    t = _WS_CACHE.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _WS_CACHE[key] = t
    return t, nbytes


def _query(name: str, *args) -> int:
    key = (name, args)
    v = _WS_QUERY_CACHE.get(key)
    if v is None:
        v = _WS_QUERY_CACHE[key] = lib().query(name, *args)
    return v


# ------------------------------------------------------------------ collator
STATUS_WORDS = 8
ST_EDGE, ST_ATOM, ST_BOND, ST_LAST_GRAPH, ST_MAX_DEGREE, ST_DEGREE_OVERFLOW = 0, 1, 2, 3, 4, 5


def csr_build(edge_index: Optional[Tensor], num_nodes: int, src32: Optional[Tensor] = None,
              dst32: Optional[Tensor] = None):
    """dst-sorted CSR + src-sorted view.  Pass either int64 ``edge_index`` [2,E] or int32 src/dst."""
    L = lib()
    if edge_index is not None:
        _chk(edge_index, torch.int64, "edge_index")
        if edge_index.dim() != 2 or edge_index.shape[0] != 2:
            raise ValueError("gnnx: edge_index must be [2, E]")
        E, dev = edge_index.shape[1], edge_index.device
    else:
        _chk(src32, torch.int32, "src"); _chk(dst32, torch.int32, "dst")
        E, dev = src32.numel(), src32.device
    i32 = dict(dtype=torch.int32, device=dev)
    rowptr, colptr = torch.empty(num_nodes + 1, **i32), torch.empty(num_nodes + 1, **i32)
    csr_src, csr_eid, pos = torch.empty(E, **i32), torch.empty(E, **i32), torch.empty(E, **i32)
    csc_pos, csc_dst = torch.empty(E, **i32), torch.empty(E, **i32)
    # status words of the batch: [0] edge endpoint out of range (written here), [1] atom feature / [2] bond feature out
    # of vocabulary (narrow_u8 / edge_codes), [3] last graph id, [4] max in-degree -- read back together (graph.py)
    err = torch.zeros(STATUS_WORDS, **i32)
    ws, nb = _ws(_query("gnnx_csr_workspace_bytes", num_nodes, E), dev)
    if edge_index is not None:
        L.call("gnnx_csr_build_i64", edge_index.data_ptr(), num_nodes, E, rowptr.data_ptr(), csr_src.data_ptr(),
               csr_eid.data_ptr(), pos.data_ptr(), colptr.data_ptr(), csc_pos.data_ptr(), csc_dst.data_ptr(),
               err.data_ptr(), _p(ws), nb, _stream())
    else:
        L.call("gnnx_csr_build_i32", src32.data_ptr(), dst32.data_ptr(), num_nodes, E, rowptr.data_ptr(),
               csr_src.data_ptr(), csr_eid.data_ptr(), pos.data_ptr(), colptr.data_ptr(), csc_pos.data_ptr(),
               csc_dst.data_ptr(), err.data_ptr(), _p(ws), nb, _stream())
    return rowptr, csr_src, csr_eid, pos, colptr, csc_pos, csc_dst, err


def degree_hist(rowptr: Tensor, hist: Tensor, max_degree: Tensor) -> None:
    _chk(rowptr, torch.int32, "rowptr"); _chk(hist, torch.int64, "hist"); _chk(max_degree, torch.int32, "max_degree")
    lib().call("gnnx_degree_hist", rowptr.data_ptr(), rowptr.numel() - 1, hist.data_ptr(), hist.numel(),
               max_degree.data_ptr(), _stream())


def graph_ptr_from_batch(batch: Tensor, num_graphs: int) -> Tensor:
    _chk(batch, torch.int64, "batch")
    gp = torch.empty(num_graphs + 1, dtype=torch.int32, device=batch.device)
    lib().call("gnnx_graph_ptr_from_batch", batch.data_ptr(), batch.numel(), num_graphs, gp.data_ptr(), _stream())
    return gp


def narrow_u8(t: Tensor, limits=None, err: Optional[Tensor] = None) -> Tensor:
    """int64 categorical features [rows, K] -> uint8.  ``limits`` (K vocabulary sizes): values outside [0, limit) are
    written as 0 and raise ``err`` (a 1-element int32 device tensor / view) -- nothing downstream can index past a table."""
    import ctypes
    _chk(t, torch.int64, "features")
    if t.dim() != 2:
        raise ValueError("gnnx.narrow_u8: features must be [rows, K]")
    rows, K = t.shape
    if limits is not None and len(limits) != K:
        raise ValueError(f"gnnx.narrow_u8: {K} feature columns but {len(limits)} vocabulary sizes")
    out = torch.empty(t.shape, dtype=torch.uint8, device=t.device)
    lim = (ctypes.c_int32 * K)(*limits) if limits is not None else None
    lib().call("gnnx_narrow_i64_u8", t.data_ptr(), rows, K, lim, out.data_ptr(), _p(err), _stream())
    return out


def edge_codes(edge_attr_u8: Tensor, csr_eid: Tensor, err: Optional[Tensor] = None) -> Tensor:
    _chk(edge_attr_u8, torch.uint8, "edge_attr_u8"); _chk(csr_eid, torch.int32, "csr_eid")
    E = csr_eid.numel()
    if edge_attr_u8.numel() != 3 * E:
        raise ValueError(f"gnnx.edge_codes: edge_attr must be [{E}, 3], got {tuple(edge_attr_u8.shape)}")
    codes = torch.empty(E, dtype=torch.uint8, device=csr_eid.device)
    lib().call("gnnx_edge_codes", edge_attr_u8.data_ptr(), csr_eid.data_ptr(), E, codes.data_ptr(), _p(err), _stream())
    return codes


def collate_packed(store, graph_ids: Tensor, node_off: Tensor, edge_off: Tensor, N: int, E: int):
    """``store`` is a ``DeviceGraphStore``; returns x_u8 [N,9], src32, dst32, edge_attr_u8 [E,3], batch32 [N]."""
    dev = graph_ids.device
    _chk(graph_ids, torch.int64, "graph_ids"); _chk(node_off, torch.int64, "node_off"); _chk(edge_off, torch.int64, "edge_off")
    G = graph_ids.numel()
    x = torch.empty((N, 9), dtype=torch.uint8, device=dev)
    src, dst = torch.empty(E, dtype=torch.int32, device=dev), torch.empty(E, dtype=torch.int32, device=dev)
    ea = torch.empty((E, 3), dtype=torch.uint8, device=dev)
    batch = torch.empty(N, dtype=torch.int32, device=dev)
    lib().call("gnnx_collate_packed", store.x.data_ptr(), store.src.data_ptr(), store.dst.data_ptr(),
               store.edge_attr.data_ptr(), store.node_ptr.data_ptr(), store.edge_ptr.data_ptr(), graph_ids.data_ptr(), G,
               node_off.data_ptr(), edge_off.data_ptr(), x.data_ptr(), src.data_ptr(), dst.data_ptr(), ea.data_ptr(),
               batch.data_ptr(), _stream())
    return x, src, dst, ea, batch


# ------------------------------------------------------------------ embeddings
def embed_sum_fwd(idx_u8: Tensor, offsets: Tensor, table: Tensor, dropout_p: float = 0.0, seed: int = 0) -> Tensor:
    """``dropout_p`` > 0: the reference's node-feature Dropout in front of the first conv, fused (see include/gnnx.h)."""
    _chk(idx_u8, torch.uint8, "idx"); _chk(offsets, torch.int32, "offsets"); _chk(table, torch.float32, "table")
    rows, K = idx_u8.shape
    H = table.shape[1]
    out = torch.empty((rows, H), dtype=torch.float32, device=table.device)
    lib().call("gnnx_embed_sum_fwd", idx_u8.data_ptr(), rows, K, offsets.data_ptr(), table.data_ptr(), H, float(dropout_p),
               int(seed), out.data_ptr(), _stream())
    return out


def embed_sum_bwd(idx_u8: Tensor, offsets: Tensor, table_rows: int, dout: Tensor, dropout_p: float = 0.0, seed: int = 0) -> Tensor:
    _chk(idx_u8, torch.uint8, "idx"); _chk(dout, torch.float32, "dout")
    rows, K = idx_u8.shape
    H = dout.shape[1]
    dtable = torch.empty((table_rows, H), dtype=torch.float32, device=dout.device)
    ws, nb = _ws(_query("gnnx_embed_sum_bwd_workspace_bytes", rows, table_rows, H), dout.device)
    lib().call("gnnx_embed_sum_bwd", idx_u8.data_ptr(), rows, K, offsets.data_ptr(), table_rows, dout.data_ptr(), H,
               float(dropout_p), int(seed), dtable.data_ptr(), _p(ws), nb, _stream())
    return dtable


def dropout_keep_mask(dropout_p: float, seed: int, shape, device) -> Tensor:
    """The node-feature dropout draw for (seed, flat element index) as a bool tensor of ``shape`` (tests replay it)."""
    keep = torch.empty(shape, dtype=torch.uint8, device=device)
    lib().call("gnnx_dropout_keep_mask", float(dropout_p), int(seed), keep.numel(), keep.data_ptr(), _stream())
    return keep.bool()


# ------------------------------------------------------------------ dense transforms
def _ld(t: Tensor) -> int:
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError("gnnx: GEMM operands must be 2-D with unit inner stride")
    return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])


def gemm(a: Tensor, b: Tensor, *, trans_a: bool = False, trans_b: bool = False, bias: Optional[Tensor] = None,
         relu: bool = False, out: Optional[Tensor] = None, accumulate: bool = False,
         precision: Optional[int] = None, relu_mask: Optional[Tensor] = None, want_relu_bits: bool = False,
         want_colsum_a: bool = False):
    """``out = act(op(a) @ op(b) + bias)``; ``a``/``b``/``out`` may be column slices (row stride = ld).
    ``relu_mask`` [M,N]: the result is zeroed where ``relu_mask <= 0`` (fused ReLU backward of the layer below); an
    int32 tensor is taken as the PACKED form (bit c%32 of word [r, c//32] set where the activation was
    positive) -- int32 [M, ceil(N/32)].  ``want_relu_bits``: also return that packed mask of the result -> ``(out, bits)``.
    ``want_colsum_a`` (``trans_a=False``): also return ``a.sum(0)`` -- the bias gradient that goes with a dX product,
    summed from the operand tiles the kernel already holds -> ``(out, colsum)``."""
    for t, n in ((a, "a"), (b, "b")):
        if not t.is_cuda or t.dtype != torch.float32:
            raise RuntimeError(f"gnnx.gemm: `{n}` must be a CUDA float32 tensor (no CPU fallback exists)")
    M, K = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
    Kb, N = (b.shape[1], b.shape[0]) if trans_b else (b.shape[0], b.shape[1])
    if K != Kb:
        raise ValueError(f"gnnx.gemm: inner dimensions differ ({K} vs {Kb})")
    if out is None:
        if accumulate:
            raise ValueError("gnnx.gemm: accumulate needs `out`")
        out = torch.empty((M, N), dtype=torch.float32, device=a.device)
    elif out.shape != (M, N) or out.dtype != torch.float32 or not out.is_cuda:
        raise ValueError("gnnx.gemm: bad `out`")
    if bias is not None:
        _chk(bias, torch.float32, "bias")
    flags = (GEMM_RELU if relu else 0) | (GEMM_ACCUMULATE if accumulate else 0)
    words = (N + 31) // 32
    ld_mask = 0
    if relu_mask is not None:
        if not relu_mask.is_cuda:
            raise ValueError("gnnx.gemm: bad `relu_mask`")
        if relu_mask.dtype == torch.int32:
            if relu_mask.shape != (M, words) or not relu_mask.is_contiguous():
                raise ValueError("gnnx.gemm: a packed `relu_mask` must be a contiguous int32 [M, ceil(N/32)]")
            flags |= GEMM_MASK_BITS
            ld_mask = words
        elif relu_mask.shape != (M, N) or relu_mask.dtype != torch.float32:
            raise ValueError("gnnx.gemm: bad `relu_mask`")
        else:
            ld_mask = _ld(relu_mask)
    bits = torch.empty((M, words), dtype=torch.int32, device=a.device) if want_relu_bits else None
    if want_colsum_a and (trans_a or want_relu_bits):
        raise ValueError("gnnx.gemm: want_colsum_a needs trans_a=False and cannot be combined with want_relu_bits")
    csum = torch.empty(K, dtype=torch.float32, device=a.device) if want_colsum_a else None
    prec = _default_precision if precision is None else precision
    if FORCE_TC and prec != PREC_FP32:
        flags |= GEMM_FORCE_TC
    L = lib()
    ws, nb = _ws(_query("gnnx_gemm_workspace_bytes", M, N, K, int(trans_a), int(trans_b), prec), a.device)
    L.call("gnnx_gemm", a.data_ptr(), _ld(a), int(trans_a), b.data_ptr(), _ld(b), int(trans_b), out.data_ptr(), _ld(out),
           M, N, K, _p(bias), flags, _p(relu_mask), ld_mask, _p(bits), _p(csum), prec, _p(ws), nb, _stream())
    if want_colsum_a:
        return out, csum
    return (out, bits) if want_relu_bits else out


def gemm_dw_batched(dys, xs, precision: Optional[int] = None):
    """``[dy.T @ x for dy, x in zip(dys, xs)]`` for products of ONE shape (dy [R, Mo], x [R, No], unit inner stride) in a
    single launch -- the weight gradients of consecutive Linear layers that see the same number of rows."""
    import ctypes
    n = len(dys)
    R, Mo, No = dys[0].shape[0], dys[0].shape[1], xs[0].shape[1]
    for dy, x in zip(dys, xs):
        if dy.shape != (R, Mo) or x.shape != (R, No) or dy.dtype != torch.float32 or x.dtype != torch.float32 or not dy.is_cuda:
            raise ValueError("gnnx.gemm_dw_batched: all products must share one shape (float32, CUDA)")
    outs = [torch.empty((Mo, No), dtype=torch.float32, device=dys[0].device) for _ in range(n)]
    prec = _default_precision if precision is None else precision
    ws, nb = _ws(_query("gnnx_gemm_dw_batched_workspace_bytes", n, Mo, No, R, prec), dys[0].device)
    vp, i64 = ctypes.c_void_p * n, ctypes.c_int64 * n
    lib().call("gnnx_gemm_dw_batched", n, vp(*[t.data_ptr() for t in dys]), i64(*[_ld(t) for t in dys]),
               vp(*[t.data_ptr() for t in xs]), i64(*[_ld(t) for t in xs]), vp(*[t.data_ptr() for t in outs]),
               i64(*[No] * n), Mo, No, R, prec, _p(ws), nb, _stream())
    return outs


def gemm_chain(x: Tensor, layers, precision: Optional[int] = None):
    """``layers`` = [(weight [W, W], bias | None, relu: bool, want_bits: bool), ...] applied in sequence to ``x`` [M, W]
    (torch Linear layout) -> ``(outs, bits)``: every layer's output and, where asked for, its packed (out > 0) mask.
    One launch on the tensor-core tier (``gnnx_gemm_chain``), the same numbers as ``gemm`` called layer by layer."""
    import ctypes
    n = len(layers)
    M, W = x.shape
    _chk(x, torch.float32, "x")
    for w, b, _, _ in layers:
        if w.shape != (W, W) or w.dtype != torch.float32 or not w.is_cuda or w.stride(1) != 1:
            raise ValueError("gnnx.gemm_chain: every weight must be a CUDA float32 [W, W] matrix with unit inner stride")
        if b is not None and (b.shape != (W,) or not b.is_contiguous()):
            raise ValueError("gnnx.gemm_chain: bias must be a contiguous [W] vector")
    if x.stride(1) != 1:
        x = x.contiguous()
    outs = [torch.empty((M, W), dtype=torch.float32, device=x.device) for _ in range(n)]
    bits = [torch.empty((M, (W + 31) // 32), dtype=torch.int32, device=x.device) if want else None for _, _, _, want in layers]
    prec = _default_precision if precision is None else precision
    vp, i64, i32 = ctypes.c_void_p * n, ctypes.c_int64 * n, ctypes.c_int32 * n
    lib().call("gnnx_gemm_chain", n, x.data_ptr(), _ld(x), vp(*[w.data_ptr() for w, _, _, _ in layers]),
               i64(*[_ld(w) for w, _, _, _ in layers]), vp(*[_p(b) for _, b, _, _ in layers]),
               i32(*[GEMM_RELU if relu else 0 for _, _, relu, _ in layers]), vp(*[o.data_ptr() for o in outs]), i64(*[W] * n),
               vp(*[_p(t) for t in bits]), M, W, prec, _stream())
    return outs, bits


def colsum(x: Tensor, out: Optional[Tensor] = None) -> Tensor:
    if not x.is_cuda or x.dtype != torch.float32:
        raise RuntimeError("gnnx.colsum: needs a CUDA float32 tensor")
    rows, cols = x.shape
    if out is None:
        out = torch.empty(cols, dtype=torch.float32, device=x.device)
    ws, nb = _ws(_query("gnnx_colsum_workspace_bytes", rows, cols), x.device)
    lib().call("gnnx_colsum", x.data_ptr(), _ld(x), rows, cols, out.data_ptr(), _p(ws), nb, _stream())
    return out


# ------------------------------------------------------------------ PNA
def pna_gather_fwd(pq: Tensor, r: Tensor, ridx: Tensor, rowptr: Tensor, csr_src: Tensor, relu: bool,
                   want_relu_bits: bool = False):
    _chk(pq, torch.float32, "pq"); _chk(r, torch.float32, "r")
    N, H = pq.shape[0], pq.shape[1] // 2
    E = csr_src.numel()
    h = torch.empty((E, H), dtype=torch.float32, device=pq.device)
    bits = torch.empty((E, (H + 31) // 32), dtype=torch.int32, device=pq.device) if want_relu_bits else None
    i32 = ridx.data_ptr() if ridx.dtype == torch.int32 else None
    u8 = ridx.data_ptr() if ridx.dtype == torch.uint8 else None
    lib().call("gnnx_pna_gather_fwd", pq.data_ptr(), pq.stride(0), r.data_ptr(), r.stride(0), i32, u8, rowptr.data_ptr(),
               csr_src.data_ptr(), N, E, H, int(relu), h.data_ptr(), _p(bits), _stream())
    return (h, bits) if want_relu_bits else h


def pna_gather_bwd(dh: Tensor, g, want_dr_edge: bool):
    _chk(dh, torch.float32, "dh")
    E, H = dh.shape
    dpq = torch.empty((g.num_nodes, 2 * H), dtype=torch.float32, device=dh.device)
    dr = torch.empty((E, H), dtype=torch.float32, device=dh.device) if want_dr_edge else None
    lib().call("gnnx_pna_gather_bwd", dh.data_ptr(), g.rowptr.data_ptr(), g.colptr.data_ptr(), g.csc_pos.data_ptr(),
               g.csr_eid.data_ptr(), g.num_nodes, E, H, dpq.data_ptr(), 2 * H, _p(dr), H, _stream())
    return dpq, dr


CODE_SORTED_MIN_ROWS = 8192      # below this the table kernel's single pass is as fast and needs no ordering
_code_order_cache: dict = {}


def code_order(codes: Tensor, num_codes: int):
    """Slots ordered by code, built ONCE per batch and shared by every layer's ``code_segment_sum``: ``(order, code_ptr,
    chunk_ptr)`` = slot ids sorted by code (stable radix sort on the uint8 keys), first position of each code, and the
    exclusive prefix sum of ceil(count / 256) (the 256-row chunks of ``gnnx_code_segment_sum_sorted`` never straddle
    two codes).  Cached on the codes tensor's storage and version (one entry: a step has one edge-code tensor)."""
    key = (codes.data_ptr(), codes.numel(), codes._version, num_codes, codes.device)
    ent = _code_order_cache.get("last")
    if ent is not None and ent[0] == key:
        if ent[3] is not None:                               # built ahead on the side stream (code_order_ahead)
            torch.cuda.current_stream().wait_event(ent[3])
        return ent[1]
    srt = torch.sort(codes, stable=True)                     # no host sync anywhere below: the call is graph-capturable
    order = srt.indices.to(torch.int32)
    ks = torch.arange(num_codes + 1, dtype=torch.int32, device=codes.device)
    code_ptr = torch.searchsorted(srt.values.to(torch.int32), ks).to(torch.int32)        # first position with code >= k
    cnt = code_ptr[1:] - code_ptr[:-1]
    chunk_ptr = torch.cat([code_ptr[:1] * 0, torch.cumsum((cnt + 255) // 256, 0).to(torch.int32)])
    res = (order, code_ptr, chunk_ptr)
    _code_order_cache["last"] = (key, res, codes, None)     # keeps `codes` alive so that the address cannot be reused
    return res


def code_order_ahead(codes: Tensor, num_codes: int) -> None:
    """Build the batch's code ordering NOW, on the side stream: only the backward needs it (the gradient of the edge-term
    tables), so its ~8 small kernels run under the forward instead of in front of the first backward layer.  No-op for
    batches below ``CODE_SORTED_MIN_ROWS`` and while a CUDA graph is being captured (the backward then builds it inline)."""
    from . import streams
    if codes.numel() < CODE_SORTED_MIN_ROWS or torch.cuda.is_current_stream_capturing():
        return
    sd = streams.side(codes.device)
    if sd is None:
        code_order(codes, num_codes)
        return
    sd.begin()
    try:
        code_order(codes, num_codes)
        ev = sd.mark()
    finally:
        sd.end()
    key, res, keep, _ = _code_order_cache["last"]
    _code_order_cache["last"] = (key, res, keep, ev)


def code_segment_sum(dh: Tensor, codes: Tensor, num_codes: int) -> Tensor:
    _chk(dh, torch.float32, "dh"); _chk(codes, torch.uint8, "codes")
    rows, H = dh.shape
    out = torch.empty((num_codes, H), dtype=torch.float32, device=dh.device)
    if rows >= CODE_SORTED_MIN_ROWS and H % 4 == 0 and dh.is_contiguous():
        order, code_ptr, chunk_ptr = code_order(codes, num_codes)
        ws, nb = _ws(_query("gnnx_code_segment_sum_sorted_workspace_bytes", rows, num_codes, H), dh.device)
        lib().call("gnnx_code_segment_sum_sorted", dh.data_ptr(), order.data_ptr(), code_ptr.data_ptr(), chunk_ptr.data_ptr(),
                   rows, num_codes, H, out.data_ptr(), _p(ws), nb, _stream())
        return out
    ws, nb = _ws(_query("gnnx_code_segment_sum_workspace_bytes", rows, num_codes, H), dh.device)
    lib().call("gnnx_code_segment_sum", dh.data_ptr(), codes.data_ptr(), rows, num_codes, H, out.data_ptr(), _p(ws), nb, _stream())
    return out


def pna_aggregate_fwd(m: Tensor, rowptr: Tensor, towers: int) -> Tensor:
    _chk(m, torch.float32, "m"); _chk(rowptr, torch.int32, "rowptr")
    N, H = rowptr.numel() - 1, m.shape[1]
    agg = torch.empty((N, 4 * H), dtype=torch.float32, device=m.device)
    lib().call("gnnx_pna_aggregate_fwd", m.data_ptr(), rowptr.data_ptr(), N, towers, H // towers, agg.data_ptr(), _stream())
    return agg


def pna_aggregate_bwd(dagg: Tensor, m: Tensor, agg: Tensor, rowptr: Tensor, towers: int) -> Tensor:
    _chk(dagg, torch.float32, "dagg"); _chk(m, torch.float32, "m"); _chk(agg, torch.float32, "agg")
    N, H = rowptr.numel() - 1, m.shape[1]
    dm = torch.empty_like(m)
    lib().call("gnnx_pna_aggregate_bwd", dagg.data_ptr(), m.data_ptr(), agg.data_ptr(), rowptr.data_ptr(), N, towers,
               H // towers, dm.data_ptr(), _stream())
    return dm


def pna_post_combine_fwd(z: Tensor, bias: Optional[Tensor], rowptr: Tensor, avg_deg_log: float, relu: bool) -> Tensor:
    """z [N, 3*Ho] = (identity | amplification | attenuation) partial products -> out [N, Ho]."""
    _chk(z, torch.float32, "z")
    N, H = z.shape[0], z.shape[1] // 3
    out = torch.empty((N, H), dtype=torch.float32, device=z.device)
    lib().call("gnnx_pna_post_combine_fwd", z.data_ptr(), _p(bias), rowptr.data_ptr(), float(avg_deg_log), N, H, int(relu),
               out.data_ptr(), _stream())
    return out


def pna_post_combine_bwd(dout: Tensor, out: Tensor, rowptr: Tensor, avg_deg_log: float, relu: bool) -> Tensor:
    _chk(dout, torch.float32, "dout"); _chk(out, torch.float32, "out")
    N, H = dout.shape
    dz = torch.empty((N, 3 * H), dtype=torch.float32, device=dout.device)
    lib().call("gnnx_pna_post_combine_bwd", dout.data_ptr(), out.data_ptr(), rowptr.data_ptr(), float(avg_deg_log), N, H,
               int(relu), dz.data_ptr(), _stream())
    return dz


# ------------------------------------------------------------------ degree-grouped first post Linear (PNAConv)
MAX_GROUPS = 16
GROUP_ALIGN = 256       # bucket padding: one 2-CTA row block of the tensor-core kernel


def degree_buckets(rowptr: Tensor, n_groups: int, err: Optional[Tensor] = None):
    """Stable bucketing of the nodes by in-degree -> ``(node_pos[N], row_node[cap], bucket_ptr[n_groups+1], err, cap)``.
    ``err`` (int32 [1], optional): where the "a node has in-degree >= n_groups" flag goes (e.g. a batch status word)."""
    _chk(rowptr, torch.int32, "rowptr")
    if not 1 <= n_groups <= MAX_GROUPS:
        raise ValueError(f"gnnx.degree_buckets: 1..{MAX_GROUPS} groups")
    N, dev = rowptr.numel() - 1, rowptr.device
    cap = _query("gnnx_degree_buckets_capacity", N, n_groups, GROUP_ALIGN)
    i32 = dict(dtype=torch.int32, device=dev)
    node_pos, row_node = torch.empty(N, **i32), torch.empty(cap, **i32)
    bucket_ptr = torch.empty(n_groups + 1, **i32)
    if err is None:
        err = torch.empty(1, **i32)
    ws, nb = _ws(_query("gnnx_degree_buckets_workspace_bytes", N, n_groups), dev)
    lib().call("gnnx_degree_buckets", rowptr.data_ptr(), N, n_groups, GROUP_ALIGN, node_pos.data_ptr(), row_node.data_ptr(),
               cap, bucket_ptr.data_ptr(), err.data_ptr(), _p(ws), nb, _stream())
    return node_pos, row_node, bucket_ptr, err, cap


def group_split_table(bucket_ptr: Tensor, n_groups: int, splits: int) -> Tensor:
    table = torch.empty((splits, 3), dtype=torch.int32, device=bucket_ptr.device)
    lib().call("gnnx_group_split_table", bucket_ptr.data_ptr(), n_groups, 32, splits, table.data_ptr(), _stream())
    return table


def pna_aggregate_rows_fwd(m: Tensor, x: Tensor, g, buckets, towers: int) -> Tensor:
    """``[agg | x]`` in degree-bucket order: [cap, 5H], padding rows zero."""
    _chk(m, torch.float32, "m"); _chk(x, torch.float32, "x")
    H = m.shape[1]
    out = torch.empty((buckets.capacity, 5 * H), dtype=torch.float32, device=m.device)
    lib().call("gnnx_pna_aggregate_rows_fwd", m.data_ptr(), g.rowptr.data_ptr(), g.num_nodes, towers, H // towers,
               buckets.row_node.data_ptr(), buckets.capacity, x.data_ptr(), out.data_ptr(), 5 * H, _stream())
    return out


def pna_post_fused_ok(buckets, towers: int, f: int, h_out: int, precision: Optional[int] = None) -> bool:
    """Whether ``pna_post_fused_fwd`` replaces ``pna_aggregate_rows_fwd`` + ``gemm_grouped`` for this layer shape."""
    prec = _default_precision if precision is None else precision
    if prec == PREC_FP32 or not FUSED_POST:
        return False
    level = _query("gnnx_pna_post_fused_supported", buckets.capacity, towers, f, h_out, prec)
    return level >= (1 if FORCE_TC else 2)


def pna_post_fused_fwd(m: Tensor, x: Tensor, g, buckets, towers: int, weff: Tensor, bias: Optional[Tensor], relu: bool,
                       want_relu_bits: bool = False, want_agg: bool = False, precision: Optional[int] = None):
    """``act([agg | x] W_eff(d)^T + b)`` in bucket order without materialising ``[agg | x]`` on the way in (one tcgen05
    kernel whose A operand is produced in shared memory from the messages) -> ``(out [cap, Ho], bits | None, agg | None)``;
    ``want_agg``: also keep the operand ([cap, 5H], what ``pna_aggregate_rows_fwd`` returns) for the backward."""
    _chk(m, torch.float32, "m"); _chk(x, torch.float32, "x"); _chk(weff, torch.float32, "weff")
    H = x.shape[1]
    G = buckets.n_groups
    Ho = weff.shape[0] // G
    cap = buckets.capacity
    if weff.shape != (G * Ho, 5 * H) or m.shape[1] != H:
        raise ValueError("gnnx.pna_post_fused_fwd: operand shapes do not match the bucket layout")
    out = torch.empty((cap, Ho), dtype=torch.float32, device=x.device)
    bits = torch.empty((cap, (Ho + 31) // 32), dtype=torch.int32, device=x.device) if want_relu_bits else None
    agg = torch.empty((cap, 5 * H), dtype=torch.float32, device=x.device) if want_agg else None
    prec = _default_precision if precision is None else precision
    lib().call("gnnx_pna_post_fused_fwd", _p(m), g.rowptr.data_ptr(), buckets.row_node.data_ptr(), cap, x.data_ptr(), towers,
               H // towers, weff.data_ptr(), _ld(weff), buckets.bucket_ptr.data_ptr(), G, _p(bias), int(relu), out.data_ptr(), Ho,
               Ho, _p(bits), _p(agg), 5 * H, prec, _stream())
    return out, bits, agg


def pna_aggregate_rows_bwd(dax: Tensor, m: Tensor, ax: Tensor, g, buckets, towers: int):
    """-> ``(dm [E,H], dx [N,H])`` from the gradient of the bucket-ordered ``[agg | x]`` operand."""
    _chk(dax, torch.float32, "dax"); _chk(m, torch.float32, "m"); _chk(ax, torch.float32, "ax")
    H = m.shape[1]
    dm = torch.empty_like(m)
    dx = torch.empty((g.num_nodes, H), dtype=torch.float32, device=m.device)
    lib().call("gnnx_pna_aggregate_rows_bwd", dax.data_ptr(), m.data_ptr(), ax.data_ptr(), 5 * H, buckets.node_pos.data_ptr(),
               g.rowptr.data_ptr(), g.num_nodes, towers, H // towers, dm.data_ptr(), dx.data_ptr(), _stream())
    return dm, dx


def pna_degree_weights(wz: Tensor, wx: Tensor, avg_deg_log: float, n_groups: int):
    """-> ``(weff [D*Ho, 4H+H], coef [D,3])`` (see include/gnnx.h)."""
    _chk(wz, torch.float32, "wz"); _chk(wx, torch.float32, "wx")
    Ho, H = wx.shape
    K4 = wz.shape[1]
    weff = torch.empty((n_groups * Ho, K4 + H), dtype=torch.float32, device=wz.device)
    coef = torch.empty((n_groups, 3), dtype=torch.float32, device=wz.device)
    lib().call("gnnx_pna_degree_weights", wz.data_ptr(), wx.data_ptr(), Ho, K4, H, float(avg_deg_log), n_groups,
               weff.data_ptr(), coef.data_ptr(), _stream())
    return weff, coef


def gather_rows(src: Tensor, idx: Tensor, cols: Optional[int] = None, want_relu_bits: bool = False):
    """``out[r] = src[idx[r], :cols]`` (zero rows where ``idx[r] < 0``); ``src`` may be a column slice."""
    _chk(idx, torch.int32, "idx")
    if not src.is_cuda or src.dtype != torch.float32:
        raise RuntimeError("gnnx.gather_rows: needs a CUDA float32 tensor")
    cols = src.shape[1] if cols is None else cols
    rows = idx.numel()
    out = torch.empty((rows, cols), dtype=torch.float32, device=src.device)
    bits = torch.empty((rows, (cols + 31) // 32), dtype=torch.int32, device=src.device) if want_relu_bits else None
    lib().call("gnnx_gather_rows", src.data_ptr(), _ld(src), idx.data_ptr(), rows, cols, out.data_ptr(), cols, _p(bits), _stream())
    return (out, bits) if want_relu_bits else out


def gemm_grouped(a: Tensor, b: Tensor, buckets, *, trans_b: bool, bias: Optional[Tensor] = None, relu: bool = False,
                 precision: Optional[int] = None, want_colsum_a: bool = False, want_relu_bits: bool = False):
    """``out[rows of bucket g] = act(a @ op(b_g) + bias)``; ``b`` stacks one weight block per group along its rows."""
    for t, n in ((a, "a"), (b, "b")):
        if not t.is_cuda or t.dtype != torch.float32:
            raise RuntimeError(f"gnnx.gemm_grouped: `{n}` must be a CUDA float32 tensor (no CPU fallback exists)")
    G = buckets.n_groups
    M, K = a.shape
    if trans_b:
        N = b.shape[0] // G
        ok = b.shape == (G * N, K)
    else:
        N = b.shape[1]
        ok = b.shape == (G * K, N)
    if not ok or M != buckets.capacity:
        raise ValueError("gnnx.gemm_grouped: operand shapes do not match the bucket layout")
    out = torch.empty((M, N), dtype=torch.float32, device=a.device)
    csum = torch.empty(K, dtype=torch.float32, device=a.device) if want_colsum_a else None
    bits = torch.empty((M, (N + 31) // 32), dtype=torch.int32, device=a.device) if want_relu_bits else None
    prec = _default_precision if precision is None else precision
    ws, nb = _ws(_query("gnnx_gemm_workspace_bytes", M, N, K, 0, int(trans_b), prec), a.device)
    lib().call("gnnx_gemm_grouped", a.data_ptr(), _ld(a), b.data_ptr(), _ld(b), int(trans_b), out.data_ptr(), N, M, N, K,
               buckets.bucket_ptr.data_ptr(), G, _p(bias), GEMM_RELU if relu else 0, None, 0, _p(bits), _p(csum), prec,
               _p(ws), nb, _stream())
    if want_relu_bits:
        return out, bits
    return (out, csum) if want_colsum_a else out


def gemm_grouped_dw(a: Tensor, b: Tensor, buckets, coef: Tensor, precision: Optional[int] = None) -> Tensor:
    """``out[j] = sum_g coef[g, j] * a[rows_g]^T @ b[rows_g]`` -> [J, Mo, No]."""
    _chk(a, torch.float32, "a"); _chk(b, torch.float32, "b"); _chk(coef, torch.float32, "coef")
    R, Mo = a.shape
    No = b.shape[1]
    J = coef.shape[1]
    prec = _default_precision if precision is None else precision
    S = _query("gnnx_gemm_grouped_dw_splits", Mo, No, R, buckets.n_groups, prec)
    table = buckets.split_table(S)
    out = torch.empty((J, Mo, No), dtype=torch.float32, device=a.device)
    ws, nb = _ws(_query("gnnx_gemm_grouped_dw_workspace_bytes", Mo, No, S), a.device)
    lib().call("gnnx_gemm_grouped_dw", a.data_ptr(), Mo, b.data_ptr(), No, R, Mo, No, table.data_ptr(), S, coef.data_ptr(), J,
               out.data_ptr(), prec, _p(ws), nb, _stream())
    return out


def relu_bwd(dy: Tensor, y: Tensor) -> Tensor:
    _chk(dy, torch.float32, "dy"); _chk(y, torch.float32, "y")
    dx = torch.empty_like(dy)
    lib().call("gnnx_relu_bwd", dy.data_ptr(), y.data_ptr(), dy.numel(), dx.data_ptr(), _stream())
    return dx


# ------------------------------------------------------------------ batch norm / pooling
def bn_relu_fwd(x: Tensor, weight, bias, running_mean, running_var, training: bool, momentum: float, eps: float,
                relu: bool, buckets=None, dropout_p: float = 0.0, seed: int = 0):
    """``buckets`` (graph.DegreeBuckets): ``x`` is a PNAConv output left in degree-bucket order [capacity, C]; the
    result is in node order [N, C] and padding rows take no part in the statistics.  ``dropout_p`` > 0: the next
    layer's node-feature Dropout applied in the same pass (draw = f(seed, node-order element index))."""
    _chk(x, torch.float32, "x")
    rows, C = x.shape
    mean = torch.empty(C, dtype=torch.float32, device=x.device)
    rstd = torch.empty(C, dtype=torch.float32, device=x.device)
    ws, nb = _ws(_query("gnnx_bn_workspace_bytes", rows, C), x.device)
    if buckets is None:
        y = torch.empty_like(x)
        lib().call("gnnx_bn_relu_fwd", x.data_ptr(), rows, C, _p(weight), _p(bias), _p(running_mean), _p(running_var),
                   int(training), float(momentum), float(eps), int(relu), float(dropout_p), int(seed), y.data_ptr(),
                   mean.data_ptr(), rstd.data_ptr(), _p(ws), nb, _stream())
    else:
        n = buckets.node_pos.numel()
        if rows != buckets.capacity:
            raise ValueError("gnnx.bn_relu_fwd: a bucket-ordered input must have `capacity` rows")
        y = torch.empty((n, C), dtype=torch.float32, device=x.device)
        lib().call("gnnx_bn_relu_rows_fwd", x.data_ptr(), rows, buckets.row_node.data_ptr(), n, C, _p(weight), _p(bias),
                   _p(running_mean), _p(running_var), int(training), float(momentum), float(eps), int(relu), float(dropout_p),
                   int(seed), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), _p(ws), nb, _stream())
    return y, mean, rstd


def bn_relu_bwd(dy: Tensor, x: Tensor, y: Optional[Tensor], weight, mean: Tensor, rstd: Tensor, relu: bool, buckets=None,
                bias=None, training: bool = True, dropout_p: float = 0.0, seed: int = 0):
    """``bias``: the forward's bias -- the ReLU mask is re-derived from ``x`` (``y`` is not read and may be ``None``).
    ``training=False``: eval-mode backward (the statistics are constants).  ``dropout_p`` / ``seed``: as in the forward."""
    _chk(dy, torch.float32, "dy")
    rows, C = x.shape
    dx = torch.empty_like(x)
    dw = torch.empty(C, dtype=torch.float32, device=x.device)
    db = torch.empty(C, dtype=torch.float32, device=x.device)
    ws, nb = _ws(_query("gnnx_bn_workspace_bytes", rows, C), x.device)
    if buckets is None:
        lib().call("gnnx_bn_relu_bwd", dy.data_ptr(), x.data_ptr(), _p(y), rows, C, _p(weight), _p(bias), mean.data_ptr(),
                   rstd.data_ptr(), int(relu), int(training), float(dropout_p), int(seed), dx.data_ptr(), dw.data_ptr(),
                   db.data_ptr(), _p(ws), nb, _stream())
    else:
        lib().call("gnnx_bn_relu_rows_bwd", dy.data_ptr(), x.data_ptr(), _p(y), rows, buckets.row_node.data_ptr(),
                   buckets.node_pos.numel(), C, _p(weight), _p(bias), mean.data_ptr(), rstd.data_ptr(), int(relu),
                   int(training), float(dropout_p), int(seed), dx.data_ptr(), dw.data_ptr(), db.data_ptr(), _p(ws), nb, _stream())
    return dx, dw, db


POOL_MODES = {"add": 0, "sum": 0, "mean": 1, "max": 2}


def segment_pool_fwd(x: Tensor, graph_ptr: Tensor, mode: int) -> Tensor:
    _chk(x, torch.float32, "x"); _chk(graph_ptr, torch.int32, "graph_ptr")
    G, C = graph_ptr.numel() - 1, x.shape[1]
    out = torch.empty((G, C), dtype=torch.float32, device=x.device)
    lib().call("gnnx_segment_pool_fwd", x.data_ptr(), graph_ptr.data_ptr(), G, C, mode, out.data_ptr(), _stream())
    return out


def segment_pool_bwd(dout: Tensor, x: Tensor, out: Tensor, graph_ptr: Tensor, mode: int) -> Tensor:
    _chk(dout, torch.float32, "dout")
    G, C = graph_ptr.numel() - 1, x.shape[1]
    dx = torch.zeros_like(x) if G == 0 else torch.empty_like(x)
    lib().call("gnnx_segment_pool_bwd", dout.data_ptr(), x.data_ptr(), out.data_ptr(), graph_ptr.data_ptr(), G, C, mode,
               dx.data_ptr(), _stream())
    return dx


# ------------------------------------------------------------------ GATv2
def gatv2_fwd(xlr: Tensor, e: Tensor, eidx: Tensor, att: Tensor, bias: Optional[Tensor], g, heads: int, c: int,
              slope: float, self_loops: bool, dropout_p: float = 0.0, seed: int = 0):
    _chk(xlr, torch.float32, "xlr"); _chk(e, torch.float32, "e"); _chk(att, torch.float32, "att")
    N, E, H = g.num_nodes, g.num_edges, heads * c
    out = torch.empty((N, H), dtype=torch.float32, device=xlr.device)
    alpha = torch.empty((E + N, heads), dtype=torch.float32, device=xlr.device)
    i32 = eidx.data_ptr() if eidx.dtype == torch.int32 else None
    u8 = eidx.data_ptr() if eidx.dtype == torch.uint8 else None
    lib().call("gnnx_gatv2_fwd", xlr.data_ptr(), e.data_ptr(), e.stride(0), i32, u8, att.data_ptr(), _p(bias),
               g.rowptr.data_ptr(), g.csr_src.data_ptr(), N, E, heads, c, float(slope), int(self_loops), float(dropout_p),
               int(seed), out.data_ptr(), alpha.data_ptr(), _stream())
    return out, alpha


def gatv2_bwd(dout: Tensor, out: Tensor, bias: Optional[Tensor], xlr: Tensor, e: Tensor, eidx: Tensor, att: Tensor,
              alpha: Tensor, g, heads: int, c: int, slope: float, self_loops: bool, dropout_p: float = 0.0,
              want_de: Optional[str] = None):
    """``want_de``: ``"slot"`` / ``"edge"`` also returns the gradient at the edge term of every edge, ``[E, H]`` in
    dst-sorted slot order / original edge order (written by the same pass: no ``gatv2_edge_grad`` call) -> 4-tuple."""
    _chk(dout, torch.float32, "dout")
    N, E, H = g.num_nodes, g.num_edges, heads * c
    dev = dout.device
    dxlr = torch.empty((N, 2 * H), dtype=torch.float32, device=dev)
    datt = torch.empty(H, dtype=torch.float32, device=dev)
    ds = torch.empty((E + N, H), dtype=torch.float32, device=dev)
    i32 = eidx.data_ptr() if eidx.dtype == torch.int32 else None
    u8 = eidx.data_ptr() if eidx.dtype == torch.uint8 else None
    if want_de not in (None, "slot", "edge"):
        raise ValueError("gnnx.gatv2_bwd: want_de must be None, 'slot' or 'edge'")
    de = torch.empty((E, H), dtype=torch.float32, device=dev) if want_de else None
    ws, nb = _ws(_query("gnnx_gatv2_bwd_workspace_bytes", N, E, heads, c), dev)
    lib().call("gnnx_gatv2_bwd", dout.data_ptr(), out.data_ptr(), _p(bias), xlr.data_ptr(), e.data_ptr(), e.stride(0), i32, u8,
               att.data_ptr(),
               alpha.data_ptr(), g.rowptr.data_ptr(), g.csr_src.data_ptr(), g.colptr.data_ptr(), g.csc_pos.data_ptr(),
               g.csc_dst.data_ptr(), N, E, heads, c, float(slope), int(self_loops), float(dropout_p), dxlr.data_ptr(),
               datt.data_ptr(), ds.data_ptr(), _p(de), g.csr_eid.data_ptr() if want_de == "edge" else None, _p(ws), nb, _stream())
    return (dxlr, datt, ds, de) if want_de else (dxlr, datt, ds)


def gatv2_edge_grad(ds: Tensor, g, hidden: int, self_loops: bool, original_order: bool) -> Tensor:
    E = g.num_edges
    de = torch.zeros((E, hidden), dtype=torch.float32, device=ds.device)
    lib().call("gnnx_gatv2_edge_grad", ds.data_ptr(), g.rowptr.data_ptr(), g.csr_src.data_ptr(),
               g.csr_eid.data_ptr() if original_order else None, g.num_nodes, E, hidden, int(self_loops), de.data_ptr(),
               _stream())
    return de


# ------------------------------------------------------------------ TransformerConv
def _eidx_ptrs(eidx: Tensor):
    if eidx.dtype == torch.int32:
        return eidx.data_ptr(), None
    if eidx.dtype == torch.uint8:
        return None, eidx.data_ptr()
    raise ValueError("gnnx: edge-term index must be int32 (csr_eid) or uint8 (bond codes)")


def dropout_seed() -> int:
    """A fresh 63-bit seed for an in-kernel dropout draw, taken from torch's CPU generator (so ``torch.manual_seed``
    makes runs repeatable) without touching the device."""
    return int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())


def tconv_fwd(qkv: Tensor, e: Tensor, eidx: Tensor, g, heads: int, c: int, skip_out: Tensor, dropout_p: float = 0.0,
              seed: int = 0):
    """``skip_out`` holds lin_skip(x) on entry and ``agg + lin_skip(x)`` on return; returns (agg, alpha)."""
    _chk(qkv, torch.float32, "qkv"); _chk(e, torch.float32, "e"); _chk(skip_out, torch.float32, "skip_out")
    N, E, H = g.num_nodes, g.num_edges, heads * c
    agg = torch.empty((N, H), dtype=torch.float32, device=qkv.device)
    alpha = torch.empty((E, heads), dtype=torch.float32, device=qkv.device)
    i32, u8 = _eidx_ptrs(eidx)
    lib().call("gnnx_tconv_fwd", qkv.data_ptr(), e.data_ptr(), e.stride(0), i32, u8, g.rowptr.data_ptr(),
               g.csr_src.data_ptr(), N, E, heads, c, float(dropout_p), int(seed), skip_out.data_ptr(), agg.data_ptr(),
               skip_out.data_ptr(), alpha.data_ptr(), _stream())
    return agg, alpha


def tconv_bwd(dout: Tensor, qkv: Tensor, e: Tensor, eidx: Tensor, alpha: Tensor, agg: Tensor, g, heads: int, c: int,
              original_order: bool, dropout_p: float = 0.0):
    """Returns (dqkv [N,3H], de [E,H]); de rows in original edge order if ``original_order`` else dst-sorted."""
    _chk(dout, torch.float32, "dout")
    N, E, H = g.num_nodes, g.num_edges, heads * c
    dev = dout.device
    dqkv = torch.empty((N, 3 * H), dtype=torch.float32, device=dev)
    ds = torch.empty((E, H), dtype=torch.float32, device=dev)
    de = torch.empty((E, H), dtype=torch.float32, device=dev)
    i32, u8 = _eidx_ptrs(eidx)
    lib().call("gnnx_tconv_bwd", dout.data_ptr(), qkv.data_ptr(), e.data_ptr(), e.stride(0), i32, u8, alpha.data_ptr(),
               agg.data_ptr(), g.rowptr.data_ptr(), g.csr_src.data_ptr(), g.csr_eid.data_ptr() if original_order else None,
               g.colptr.data_ptr(), g.csc_pos.data_ptr(), g.csc_dst.data_ptr(), N, E, heads, c, float(dropout_p),
               dqkv.data_ptr(), ds.data_ptr(), de.data_ptr(), _stream())
    return dqkv, de


def assemble_params(msigmae: Tensor, munanb: Tensor, mw: Tensor, log10assoc: Optional[Tensor] = None,
                    assoc: Optional[Tensor] = None) -> Tensor:
    """``[G, 9]`` = [m, sigma, epsilon/k, kappa_AB, epsilon_AB/k, mu, n_a, n_b, MW] (demo/utils.py:664-691)."""
    G = msigmae.shape[0]
    for t, shape, n in ((msigmae, (G, 3), "msigmae"), (munanb, (G, 3), "munanb"), (mw.reshape(-1), (G,), "mw")):
        _chk(t, torch.float32, n)
        if tuple(t.shape) != shape:
            raise ValueError(f"gnnx.assemble_params: `{n}` must be {shape}, got {tuple(t.shape)}")
    for t, n in ((log10assoc, "log10assoc"), (assoc, "assoc")):
        if t is not None:
            _chk(t, torch.float32, n)
            if tuple(t.shape) != (G, 2):
                raise ValueError(f"gnnx.assemble_params: `{n}` must be ({G}, 2)")
    out = torch.empty((G, 9), dtype=torch.float32, device=msigmae.device)
    lib().call("gnnx_assemble_params", msigmae.data_ptr(), _p(log10assoc), _p(assoc), munanb.data_ptr(),
               mw.reshape(-1).data_ptr(), G, out.data_ptr(), _stream())
    return out


# ------------------------------------------------------------------ loss / optimizer tail
def relhuber_fwd_bwd(pred: Tensor, target: Tensor, delta: float = 0.01, want_grad: bool = True):
    _chk(pred, torch.float32, "pred"); _chk(target, torch.float32, "target")
    out2 = torch.empty(2, dtype=torch.float32, device=pred.device)
    dpred = torch.empty_like(pred) if want_grad else None
    lib().call("gnnx_relhuber_fwd_bwd", pred.data_ptr(), target.data_ptr(), pred.numel(), float(delta), out2.data_ptr(),
               _p(dpred), _stream())
    return out2, dpred


_param_epoch = 0


def param_epoch() -> int:
    """Bumped whenever a gnnx kernel rewrites parameters in place (the optimizer step works on raw pointers, so
    ``Tensor._version`` does not move); caches of parameter-derived tensors key on it."""
    return _param_epoch


def adamw_amsgrad_(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, max_exp_avg_sq: Tensor, *, lr: float,
                   beta1: float, beta2: float, eps: float, weight_decay: float, step: int, grad_scale: float = 1.0) -> None:
    for t, n in ((param, "param"), (grad, "grad"), (exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq"), (max_exp_avg_sq, "max_exp_avg_sq")):
        _chk(t, torch.float32, n)
    lib().call("gnnx_adamw_amsgrad", param.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
               max_exp_avg_sq.data_ptr(), param.numel(), float(lr), float(beta1), float(beta2), float(eps),
               float(weight_decay), int(step), float(grad_scale), _stream())
    global _param_epoch
    _param_epoch += 1
