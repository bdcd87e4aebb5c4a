"""The C-ABI shared library: it loads without a GPU, exports every symbol include/insb200.h declares, the
ctypes binding covers exactly that set, and creating a context without a CUDA device fails loudly (no CPU
fallback).  No compute call is made here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "insb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(insb200_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    import insb200  # noqa: F401  (builds nothing; requires the in-tree library)
    from insb200 import _lib

    lib = _lib.load()
    syms = declared_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/insb200.h but not exported by libinsb200.so"
    assert sorted(_lib.PROTOTYPES) == syms, set(_lib.PROTOTYPES) ^ set(syms)
    assert lib.insb200_version() >= 100


def test_create_validates_arguments_and_needs_a_gpu():
    from insb200 import _lib

    lib = _lib.load()
    h = C.c_void_p()
    n = (C.c_int * 3)(16, 16, 16)
    hh = (C.c_double * 3)(0.1, 0.1, 0.1)
    # argument validation happens before any CUDA call
    assert lib.insb200_create(C.byref(h), 4, n, 0, hh, 100.0, 0, 0, 1, None) != 0
    assert b"D must be 2 or 3" in lib.insb200_last_error(None)
    odd = (C.c_int * 3)(15, 16, 16)
    assert lib.insb200_create(C.byref(h), 3, odd, 0, hh, 100.0, 0, 0, 1, None) != 0
    assert b"even number of volumes" in lib.insb200_last_error(None)
    assert lib.insb200_create(C.byref(h), 3, n, 0, hh, 100.0, 0, 0, 3, None) != 0  # 16 % 3
    assert b"divisible" in lib.insb200_last_error(None)
    import torch

    if not torch.cuda.is_available():
        assert lib.insb200_create(C.byref(h), 3, n, 0, hh, 100.0, 0, 0, 1, None) != 0
        assert b"no CPU fallback" in lib.insb200_last_error(None)
        assert not h.value


def test_python_layer_has_no_fallback():
    """The host mirror holds no arithmetic of the path: without the library it cannot be imported."""
    from insb200 import _lib

    assert os.path.exists(_lib.LIB_PATH)
    src = open(os.path.join(ROOT, "incompressiblenavierstokes.jl_b200", "api.py")).read()
    assert "oracle" not in src
