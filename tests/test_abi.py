"""The C-ABI library loads and exports every symbol include/gnnx.h declares (no compute calls)."""
import ctypes
import re

from gnnepcsaft_b200 import _lib


def test_header_parses_all_prototypes():
    text = _lib.HEADER.read_text()
    declared = set(re.findall(r"GNNX_API[^;(]*?(gnnx_\w+)\s*\(", text))
    protos = _lib.parse_header()
    assert declared == set(protos) and len(protos) >= 35


def test_library_exports_every_declared_symbol():
    assert _lib.LIB_PATH.exists(), "libgnnx_sm100.so missing: run python -m gnnepcsaft_b200.build"
    dll = ctypes.CDLL(str(_lib.LIB_PATH))
    for name in _lib.parse_header():
        assert hasattr(dll, name), name
    assert dll.gnnx_abi_version() == _lib.ABI_VERSION == _lib._header_abi_version()


def test_loaded_library_matches_the_sources_on_disk():
    """a stale .so with the same symbols but another argument layout must not be callable.  The build digest
    stamped into the binary equals the digest of include/gnnx.h + csrc/ + flags on disk (the loader rebuilds or raises)."""
    from gnnepcsaft_b200 import build
    L = _lib.lib()
    assert L._dll.gnnx_build_digest().decode() == build.source_digest()
    assert _lib._stale(L._dll) == ""


def test_unknown_precisions_are_rejected():
    import pytest
    from gnnepcsaft_b200 import ops
    with pytest.raises(ValueError):
        ops.set_matmul_precision("bf16")           # no bf16 kernels are built; "medium" is served by the tf32 tier
    ops.set_matmul_precision("medium")
    assert ops.get_matmul_precision() == ops.PREC_TF32
    ops.set_matmul_precision("fp32")


def test_no_torch_types_in_signatures():
    for name, (restype, argtypes) in _lib.parse_header().items():
        for a in argtypes:
            assert a in (ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_int, ctypes.c_size_t, ctypes.c_float), name


def test_cpu_tensor_is_rejected_loudly():
    import pytest
    import torch
    from gnnepcsaft_b200 import ops
    with pytest.raises(RuntimeError):
        ops.gemm(torch.zeros(2, 2), torch.zeros(2, 2))
    with pytest.raises(RuntimeError):
        ops.pna_aggregate_fwd(torch.zeros(2, 4), torch.zeros(2, dtype=torch.int32), 1)
