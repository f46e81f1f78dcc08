"""C-ABI surface (CPU only, no compute calls): libasdmatch_cuda.so loads, exports exactly the functions
include/asdmatch.h declares, and refuses to create a context without a GPU (there is no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "asdmatch.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(asdm_[a-z0-9_]+)\s*\(", src)))


def test_header_compiles_as_c():
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HEADER])


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg._lib.load()
    names = declared_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/asdmatch.h but not exported"
    assert sorted(pkg._lib.EXPORTS) == names, "python binding list and header disagree"
    out = subprocess.check_output(["nm", "-D", "--defined-only", pkg._lib.LIB_PATH], text=True)
    exported = sorted(set(re.findall(r" T (asdm_[a-z0-9_]+)$", out, flags=re.M)))
    assert exported == names, "the .so exports C symbols the header does not declare (or the reverse)"


def test_no_cpu_fallback_without_gpu(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = pkg._lib.load()
    h = C.c_void_p()
    assert lib.asdm_create(0, C.byref(h)) == -2  # ASDM_ERR_CUDA
    assert not h.value
    with pytest.raises(pkg.AsdmError):
        pkg.Context(0)


def test_product_never_imports_oracle():
    """The package and the CUDA sources must not reference oracle/ (the oracle is test infrastructure)."""
    for base, _, files in os.walk(os.path.join(ROOT, "asd-slam_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".cpp")) or f == "Makefile":
                txt = open(os.path.join(base, f), errors="ignore").read()
                assert "oracle_py" not in txt and "liboracle" not in txt and "asd_oracle" not in txt, os.path.join(base, f)


def header_prototypes():
    """{name: number of parameters} of every function prototype in include/asdmatch.h."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"typedef[^;]*\(\s*\*[^;]*;", "", src)  # function-pointer typedefs are not prototypes
    out = {}
    for name, args in re.findall(r"\b(asdm_[a-z0-9_]+)\s*\(([^()]*)\)\s*;", src, flags=re.S):
        args = args.strip()
        out[name] = 0 if args in ("", "void") else len(args.split(","))
    return out


def test_ctypes_argument_counts_match_the_header(pkg):
    """A ctypes binding with one argument too few or too many corrupts the call silently: every function whose argtypes
    the mirror sets must take as many parameters as its prototype declares."""
    lib = pkg._lib.load()
    protos = header_prototypes()
    assert set(protos) == set(declared_functions())
    checked = 0
    for name, n in protos.items():
        at = getattr(lib, name).argtypes
        if at is None:
            continue
        assert len(at) == n, f"{name}: the header declares {n} parameters, the ctypes mirror passes {len(at)}"
        checked += 1
    assert checked >= 40
