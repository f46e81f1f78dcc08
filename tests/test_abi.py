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
