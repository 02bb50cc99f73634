"""Build and load ``libfedrec_b200.so`` (the hand-written sm_100a kernels) through its C ABI.

The library is plain ``nvcc`` output with ``extern "C"`` entry points declared in
``include/fedrec_b200.h``; Python binds it with ``ctypes`` and passes raw device
pointers plus the current CUDA stream.  There is no torch C++ extension layer: the
build takes seconds, has no ABI coupling to a torch version, and the same library
can be called from C/C++.

The ``.so`` is built in-tree next to this file so it travels with the repository
snapshot to the GPU box.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(os.path.dirname(_HERE))
_CSRC = os.path.join(_REPO, "csrc")
_INCLUDE = os.path.join(_REPO, "include")
_LIB = os.path.join(_HERE, "libfedrec_b200.so")
_OBJ = os.path.join(_REPO, "build", "obj")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr",
]

_lock = threading.Lock()
_lib: Optional[ctypes.CDLL] = None
_load_error: Optional[str] = None


def native_lib_path() -> str:
    return _LIB


def sources() -> List[str]:
    return sorted(os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith(".cu"))


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def _stale() -> bool:
    if not os.path.exists(_LIB):
        return True
    t = os.path.getmtime(_LIB)
    deps = sources() + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h"))]
    deps += [os.path.join(_INCLUDE, f) for f in os.listdir(_INCLUDE)] if os.path.isdir(_INCLUDE) else []
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a and link the shared library. Returns its path."""
    global _lib
    with _lock:
        if not force and not _stale():
            return _LIB
        os.makedirs(_OBJ, exist_ok=True)
        nvcc = _nvcc()
        hdr_time = max([os.path.getmtime(os.path.join(_CSRC, f)) for f in os.listdir(_CSRC)
                        if f.endswith((".cuh", ".h"))] + [0.0])

        def compile_one(src: str) -> str:
            obj = os.path.join(_OBJ, os.path.basename(src)[:-3] + ".o")
            if (not force and os.path.exists(obj) and os.path.getmtime(obj) > os.path.getmtime(src)
                    and os.path.getmtime(obj) > hdr_time):
                return obj
            cmd = [nvcc] + NVCC_FLAGS + ["-I", _INCLUDE, "-I", _CSRC, "-c", src, "-o", obj]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
            return obj

        with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
            objs = list(ex.map(compile_one, sources()))
        tmp = _LIB + ".tmp"
        cmd = [nvcc, "-shared", "-o", tmp] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, _LIB)
        _lib = None
        return _LIB


def _bind(lib: ctypes.CDLL) -> None:
    from ._abi import SIGNATURES
    for name, (restype, argtypes) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:  # library older than the Python side
            raise RuntimeError(f"{_LIB} lacks symbol {name}; rebuild with fedrec_b200.ops.native.build(force=True)") from e
        fn.restype = restype
        fn.argtypes = argtypes


def lib() -> ctypes.CDLL:
    global _lib, _load_error
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB) or (os.environ.get("FEDREC_AUTOBUILD", "1") != "0" and _stale()):
        # a fresh checkout has sources only (the .so is a build artefact): build it now, ~20 s with nvcc on PATH
        try:
            build()
        except Exception as e:  # noqa: BLE001
            if not os.path.exists(_LIB):
                _load_error = f"{_LIB} is not built and building it failed: {e}"
                raise RuntimeError(_load_error) from e
    with _lock:
        if _lib is None:
            if not os.path.exists(_LIB):
                _load_error = f"{_LIB} not built; run `python -c 'import __graft_entry__ as g; g.build()'`"
                raise RuntimeError(_load_error)
            try:
                # FEDREC_LIB: load another build of the same ABI (same-box A/B of a kernel change; scripts/ab_lib.sh)
                l = ctypes.CDLL(os.environ.get("FEDREC_LIB") or _LIB, mode=ctypes.RTLD_GLOBAL)
                _bind(l)
            except OSError as e:
                _load_error = str(e)
                raise RuntimeError(f"cannot load {_LIB}: {e}") from e
            _lib = l
    return _lib


def has_native() -> bool:
    try:
        lib()
        return True
    except RuntimeError:
        return False


def require_native() -> ctypes.CDLL:
    """Return the library or raise: CUDA tensors never fall back to library kernels silently."""
    return lib()


def last_error() -> str:
    return lib().fr_last_error().decode()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise RuntimeError(f"fedrec_b200 native call {what} failed (rc={rc}): {last_error()}")
