"""The native library builds for sm_100a on a CPU-only box and contains Blackwell tensor-core code."""
import os
import shutil
import subprocess

import pytest

from fedrec_b200.ops import native
from fedrec_b200.ops._abi import SIGNATURES


def test_library_builds_and_binds():
    path = native.build()
    assert os.path.exists(path)
    lib = native.lib()
    assert lib.fr_version() >= 100
    for name in SIGNATURES:
        assert hasattr(lib, name), name


def test_header_declares_every_bound_symbol():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "fedrec_b200.h")).read()
    for name in SIGNATURES:
        assert name + "(" in hdr, f"{name} missing from include/fedrec_b200.h"


@pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="cuobjdump not installed")
def test_sass_has_tcgen05_and_tma():
    sass = subprocess.run(["cuobjdump", "-sass", native.build()], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass            # tcgen05.mma kind::f16
    assert "UTMALDG" in sass            # TMA tensor load
    assert "LDTM" in sass               # tcgen05.ld
    assert "HGMMA" not in sass and "sm_100a" in sass


def test_ctypes_signatures_match_header_arity():
    """Every bound function has as many ctypes argument types as the header declares parameters."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "fedrec_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    decls = dict()
    for m in re.finditer(r"\b(\w+)\s*\(([^;{}]*?)\)\s*;", hdr):
        name, args = m.group(1), m.group(2).strip()
        n = 0 if args in ("", "void") else args.count(",") + 1
        decls[name] = n
    for name, (_res, argtypes) in SIGNATURES.items():
        assert name in decls, name
        assert decls[name] == len(argtypes), (name, decls[name], len(argtypes))


def build_c_demo():
    """Compile and link examples/c_abi_demo.c as C99 against the header and the in-tree library; returns the binary path."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_path = native.build()
    lib_dir = os.path.dirname(lib_path)
    out_dir = os.path.join(root, "build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "c_abi_demo")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = ["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(root, "include"), "-I", os.path.join(cuda, "include"),
           os.path.join(root, "examples", "c_abi_demo.c"), "-o", exe, "-L", lib_dir, "-lfedrec_b200",
           "-L", os.path.join(cuda, "lib64"), "-lcudart", "-lm", f"-Wl,-rpath,{lib_dir}",
           f"-Wl,-rpath,{os.path.join(cuda, 'lib64')}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_header_is_valid_c_and_links_from_a_c_host():
    """The public header compiles as C99 and a C program links against the library (run on a GPU: test_gpu_small_ops)."""
    exe = build_c_demo()
    assert os.path.exists(exe)
