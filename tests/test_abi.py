"""CPU tests: the C-ABI shared library loads and exports exactly what include/gprb200.h declares."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "gprb200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gprb200_[a-z_0-9]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from gpr_b200 import capi
    lib = capi.load_library()
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/gprb200.h but not exported by libgprb200.so"
    assert sorted(capi._PROTOS) == syms, "gpr_b200/capi.py prototypes and include/gprb200.h diverged"
    assert b"sm_100a" in lib.gprb200_version()


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from gpr_b200 import Handle, GprError
    with pytest.raises(GprError) as e:
        Handle(0)
    assert "no CPU fallback" in str(e.value)


def test_product_package_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "gpr_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+[\"<][^\n]*oracle|dlopen[^\n]*oracle|CDLL[^\n]*oracle",
                                     src, flags=re.M), f"{f} reaches into oracle/"


def test_dist_library_exports_every_declared_symbol():
    """include/gprb200_dist.h (multi-GPU Learn, NCCL from C++) against libgprb200_dist.so and its ctypes prototypes."""
    from gpr_b200 import capi_dist
    lib = capi_dist.load_dist_library()
    txt = open(os.path.join(ROOT, "include", "gprb200_dist.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    syms = sorted(set(re.findall(r"\b(gprb200_dist_[a-z_0-9]+)\s*\(", txt)))
    assert len(syms) == 11
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/gprb200_dist.h but not exported by libgprb200_dist.so"
    assert sorted(capi_dist._PROTOS) == syms


def test_dist_create_rejects_bad_grids_without_a_gpu_call():
    from gpr_b200 import capi_dist
    lib = capi_dist.load_dist_library()
    import ctypes as C
    h = C.c_void_p()
    assert lib.gprb200_dist_create(C.byref(h), 0, 0, 4, b"\0" * 128, 3, 2) == -2
    assert b"Pr * Pc" in lib.gprb200_dist_last_error(None)
    assert lib.gprb200_dist_create(C.byref(h), 0, 5, 4, b"\0" * 128, 2, 2) == -2
    assert lib.gprb200_dist_create(C.byref(h), 0, 1, 2, None, 1, 2) == -2
