"""ctypes binding of ``libgnnx_sm100.so`` -- the only way the package reaches the GPU kernels.

Prototypes are parsed from ``include/gnnx.h`` so the Python side can never drift from the C ABI.
There is no CPU fallback: if the shared library is missing or a call fails, a ``RuntimeError`` is
raised (the product path must fail loudly, never silently degrade).
"""

from __future__ import annotations

import ctypes
import re
from pathlib import Path
from typing import Dict, List, Tuple

_HERE = Path(__file__).resolve().parent
HEADER = _HERE.parent / "include" / "gnnx.h"
LIB_PATH = _HERE / "libgnnx_sm100.so"

_SCALARS = {
    "int64_t": ctypes.c_int64, "int32_t": ctypes.c_int32, "int": ctypes.c_int,
    "size_t": ctypes.c_size_t, "float": ctypes.c_float,
}
_PROTO = re.compile(r"GNNX_API\s+(const\s+char\s*\*|int64_t|int32_t|int|size_t)\s+(gnnx_\w+)\s*\(([^;]*?)\)\s*;", re.S)


def parse_header(path: Path = HEADER) -> Dict[str, Tuple[object, List[object]]]:
    """Return ``{name: (restype, [argtypes])}`` for every ``GNNX_API`` prototype in the header."""
    text = re.sub(r"/\*.*?\*/", "", path.read_text(), flags=re.S)
    out: Dict[str, Tuple[object, List[object]]] = {}
    for ret, name, args in _PROTO.findall(text):
        ret = ret.replace(" ", "")
        restype = ctypes.c_char_p if ret.startswith("constchar") else _SCALARS[ret]
        argtypes: List[object] = []
        args = args.strip()
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                else:
                    ctype = a.replace("const", "").split()[0]
                    argtypes.append(_SCALARS[ctype])
        out[name] = (restype, argtypes)
    return out


class GnnxError(RuntimeError):
    pass


ABI_VERSION = 2


def _header_abi_version(path: Path = HEADER) -> int:
    m = re.search(r"#define\s+GNNX_ABI_VERSION\s+(\d+)", path.read_text())
    return int(m.group(1)) if m else -1


def _stale(dll) -> str:
    """'' when the loaded binary was built from the sources on disk, else why not.  All pointers cross the boundary as
    c_void_p, so a stale .so with the same symbol names but another argument layout would corrupt device memory
    instead of failing: the build digest (header + every csrc file + flags) is stamped into the library and compared."""
    from . import build as _b
    try:
        fn = dll.gnnx_build_digest
    except AttributeError:
        return "it predates gnnx_build_digest()"
    fn.restype, fn.argtypes = ctypes.c_char_p, []
    have, want = (fn() or b"").decode(), _b.source_digest()
    return "" if have == want else f"its build digest {have[:12]} differs from the sources' {want[:12]}"


class _Lib:
    def __init__(self):
        self.graph_replayed = 0          # gnnx kernels re-launched by CUDA-graph replays (see launch_count)
        if not LIB_PATH.exists():
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m gnnepcsaft_b200.build` "
                "(there is no CPU or PyTorch fallback for the gnnx kernels)")
        self._dll = ctypes.CDLL(str(LIB_PATH))
        why = _stale(self._dll)
        if why:
            # rebuild in place once (nvcc is part of the image); a library that still does not match is refused
            from . import build as _b
            try:
                _b.build(force=True)
            except Exception as e:  # noqa: BLE001
                raise RuntimeError(f"{LIB_PATH.name} is stale ({why}) and could not be rebuilt: {e}") from e
            # dlopen caches by path: load the fresh file through a unique name so the old mapping is not reused
            import shutil
            import tempfile
            tmp = Path(tempfile.mkdtemp(prefix="gnnx_")) / LIB_PATH.name
            shutil.copy2(LIB_PATH, tmp)
            self._dll = ctypes.CDLL(str(tmp))
            why = _stale(self._dll)
            if why:
                raise RuntimeError(f"{LIB_PATH.name} is stale after a rebuild ({why}); refusing to call it")
        self.protos = parse_header()
        for name, (restype, argtypes) in self.protos.items():
            fn = getattr(self._dll, name)          # AttributeError if the .so lacks a declared symbol
            fn.restype, fn.argtypes = restype, argtypes
        abi = self._dll.gnnx_abi_version()
        if abi != ABI_VERSION or _header_abi_version() != ABI_VERSION:
            raise RuntimeError(f"libgnnx_sm100.so ABI {abi} / header ABI {_header_abi_version()} do not match the "
                               f"Python layer ({ABI_VERSION})")

    def last_error(self) -> str:
        msg = self._dll.gnnx_last_error()
        return msg.decode() if msg else ""

    def call(self, name: str, *args):
        """Call an ``int``-returning entry point and raise on a non-zero code."""
        if self._timing is not None:
            return self._timed_call(name, args)
        rc = getattr(self._dll, name)(*args)
        if rc != 0:
            raise GnnxError(f"{name} failed with code {rc}: {self.last_error()}")

    # -- optional per-entry-point device timing (bench.py's roofline pass): CUDA events on the launching stream
    _timing = None

    def start_timing(self) -> None:
        self._timing = []

    def stop_timing(self):
        """Returns ``[(name, args, ms)]`` for every call since ``start_timing()`` (synchronises the device)."""
        import torch
        torch.cuda.synchronize()
        rec, self._timing = self._timing, None
        return [(n, a, e0.elapsed_time(e1)) for n, a, e0, e1 in rec]

    def _timed_call(self, name: str, args):
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = getattr(self._dll, name)(*args)
        e1.record()
        if rc != 0:
            raise GnnxError(f"{name} failed with code {rc}: {self.last_error()}")
        self._timing.append((name, args, e0, e1))

    def trace_begin(self) -> None:
        """Open a per-kernel device timeline (two CUDA events around every kernel launch of the library)."""
        self._dll.gnnx_trace_begin()

    def trace_end(self):
        """Close it: ``[(kernel name, kernel us, us since the previous gnnx kernel ended)]`` in launch order."""
        n = int(self._dll.gnnx_trace_end(None, 0))
        buf = ctypes.create_string_buffer(n + 16)
        self._dll.gnnx_trace_end(buf, n + 16)
        rows = []
        for line in buf.value.decode().splitlines():
            name, us, gap = line.rsplit("\t", 2)
            rows.append((name, float(us), float(gap)))
        return rows

    def launch_count(self) -> int:
        """Kernels launched by libgnnx so far in this process, those replayed from a CUDA graph of gnnx kernels included
        (``graph_replayed``: the library counts a captured launch once, at capture time; the owner of the graph adds the
        captured count at every replay)."""
        return int(self._dll.gnnx_launch_count()) + self.graph_replayed

    def tc_launch_count(self) -> int:
        """Of those, tcgen05 tensor-core kernels (k_gemm_tf32_rows2 / k_gemm_tf32_red*)."""
        return int(self._dll.gnnx_tc_launch_count())

    def query(self, name: str, *args) -> int:
        """Call a ``size_t``-returning ``*_workspace_bytes`` entry point."""
        return int(getattr(self._dll, name)(*args))


_LIB = None


def lib() -> _Lib:
    global _LIB
    if _LIB is None:
        _LIB = _Lib()
    return _LIB
