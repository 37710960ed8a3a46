"""CPU tests of the host-side mirror: kernel strings, program lowering, MatrixIO byte format."""
import glob
import os

import numpy as np
import pytest

import gpr_b200 as g
from tests.synth import KERNELS

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("f", sorted(glob.glob(os.path.join(GOLD, "*.npz"))))
def test_tostring_and_parameter_order_match_reference(f):
    z = np.load(f)
    if "tostring" not in z.files:
        pytest.skip("no kernel string in this fixture")
    k = g.KernelFactory.GetKernel(str(z["kernel"]))
    assert k.ToString() == str(z["tostring"])            # Kernel::ToString of the reference
    assert k.GetParameters() == list(z["params"])        # Kernel::GetParameters order
    k2 = g.KernelFactory.GetKernel(k.ToString())
    assert k2 == k


def test_json_style_kernel_string():  # scripts/params/pmb/config_ZL0001_vol01.yaml:49
    k = g.KernelFactory.GetKernel("GaussianKernel(35, 30,)")
    assert k.GetParameters() == [35.0, 30.0] and k.ToString() == "GaussianKernel(35,30,)"


def test_lowering_is_postorder_with_k1_first():
    k = g.KernelFactory.GetKernel(KERNELS["general13"])
    ops, params = k.lower()
    assert ops == [1, 1, 5, 7, 6, 4, 6, 1, 3, 6, 6]
    assert params == k.GetParameters() and len(params) == 13


def test_general_kernel_matches_factory_string():  # include/KernelUtils.h:45-89
    p = [1.1, 2, 0.8, 3, 0.7, 1.3, 1.1, 0.9, 2.2, 1.5, 0.5, 0.8, 0.2]
    assert g.GetGeneralKernel(p).ToString() == g.KernelFactory.GetKernel(KERNELS["general13"]).ToString()


def test_zero_parameters_throw_like_the_reference():  # tests/GaussianProcessTest.cpp:322-344
    with pytest.raises(ValueError):
        g.GaussianKernel(0)
    with pytest.raises(ValueError):
        g.GaussianKernel(1, 0)
    with pytest.raises(ValueError):
        g.PeriodicKernel(1, 0, 1)


def test_set_parameters_roundtrip():
    k = g.KernelFactory.GetKernel(KERNELS["sum3"])
    p = k.GetParameters()
    k.SetParameters([v * 1.5 for v in p])
    assert k.GetParameters() == [v * 1.5 for v in p]
    with pytest.raises(ValueError):
        k.SetParameters(p[:-1])


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_matrix_io_format(tmp_path, dtype):  # lib/MatrixIO.cpp:77-99, tests/IOTest.cpp:45-90
    m = np.random.default_rng(0).standard_normal((10, 3)).astype(dtype)
    fn = str(tmp_path / "m.txt")
    g.WriteMatrix(m, fn, dtype)
    raw = open(fn, "rb").read()
    assert raw.startswith(b"10 3\n") and len(raw) == 5 + 30 * np.dtype(dtype).itemsize
    assert raw[5:] == m.tobytes()
    assert np.array_equal(g.ReadMatrix(fn, dtype), m.astype(np.float64))
    g.WriteMatrix(np.zeros((0, 0)), fn)
    assert g.ReadMatrix(fn).shape == (0, 0)
