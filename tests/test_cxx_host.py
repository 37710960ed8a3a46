"""CPU tests of the C++ host API (gpr_b200/cxx): tests/cxx/host_test.cpp covers kernel arithmetic and gradient order,
kernel-string round trips, MatrixIO, error strings and the loud failure without a GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cxx", "bin")


def _have_gpu():
    import torch
    return torch.cuda.is_available()


def test_host_library_builds_and_exports_the_reference_classes():
    lib = os.path.join(ROOT, "gpr_b200", "cxx", "libgplib.so")
    assert os.path.exists(lib), "gpr_b200/cxx/libgplib.so missing: run __graft_entry__.build()"
    syms = subprocess.run(["nm", "-DC", "--defined-only", lib], capture_output=True, text=True).stdout
    for want in ("gpr::GaussianProcess<double>::AddSample", "gpr::GaussianProcess<double>::Initialize",
                 "gpr::GaussianProcess<double>::Predict", "gpr::GaussianProcess<double>::PredictDerivative",
                 "gpr::GaussianProcess<double>::GetCredibleInterval", "gpr::GaussianProcess<double>::Save",
                 "gpr::GaussianProcess<double>::Load", "gpr::GaussianProcess<float>::Predict",
                 "gpr::GaussianProcess<double>::ComputeDerivativeKernelMatrix", "gpr::MatrixIOTest"):
        assert want in syms, f"{want} not exported by libgplib.so"
    # the host library reaches the GPU only through the C-ABI
    undef = subprocess.run(["nm", "-D", "--undefined-only", lib], capture_output=True, text=True).stdout
    assert "gprb200_learn" in undef and "gprb200_predict" in undef
    assert "dgetrf_" not in undef and "dpotrf_" not in undef, "host library must not carry a LAPACK CPU path"


def test_host_logic_program():
    exe = os.path.join(BIN, "host_test")
    assert os.path.exists(exe), "tests/cxx/bin/host_test missing: run __graft_entry__.build()"
    env = dict(os.environ)
    if not _have_gpu():
        env["GPRB200_EXPECT_NO_GPU"] = "1"
    p = subprocess.run([exe], capture_output=True, text=True, env=env, timeout=120)
    assert p.returncode == 0 and "[failed]" not in p.stdout, p.stdout + p.stderr
