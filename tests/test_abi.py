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
