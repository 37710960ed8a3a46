"""-m gpu: the reference's deterministic known-answer tests, run through the API mirror
(gpr_b200.GaussianProcess) with the reference's own sample points, test points and thresholds (SURVEY.md section 4 /
8c). The same programs also run unmodified in C++ (oracle/Makefile: dropin_tests, tests/test_gpu_cxx.py)."""
import math

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G():
    import gpr_b200
    return gpr_b200


def _sin_gp(G, sigma_k, inv=None):
    gp = G.GaussianProcess(G.GaussianKernel(sigma_k))
    gp.SetSigma(0.0)
    if inv is not None:
        gp.SetInversionMethod(inv)
    n = 10
    for i in range(n):
        x = i * 2 * math.pi / n
        gp.AddSample([x], [math.sin(x)])
    return gp


def test_gaussianprocesstest_t1_sin_regression(G):  # tests/GaussianProcessTest.cpp:35-76
    gp = _sin_gp(G, 2.889)
    err = sum(abs(gp.Predict([i * 2 * math.pi / 50])[0] - math.sin(i * 2 * math.pi / 50)) for i in range(50))
    assert err <= 8e-4  # the reference's own threshold (tests/GaussianProcessTest.cpp:71)


@pytest.mark.parametrize("inv", [0, 1, 2, 3])
def test_inversionmethodstest_all_enum_values(G, inv):  # tests/InversionMethodsTest.cpp:36-145
    gp = _sin_gp(G, 2.8, inv)   # tests/InversionMethodsTest.cpp:37
    pts = [i * 2 * math.pi / 50 for i in range(50)]
    err = sum(abs(gp.Predict([x])[0] - math.sin(x)) for x in pts)
    assert err <= 6e-4  # tests/InversionMethodsTest.cpp:80,98,117,139


def test_gaussianprocesstest_t2_two_in_two_out(G):  # tests/GaussianProcessTest.cpp:78-121
    gp = G.GaussianProcess(G.GaussianKernel(3.24))
    gp.SetSigma(0.0)
    n = 10
    for i in range(n):
        a = i * 2 * math.pi / n
        gp.AddSample([a, a], [math.sin(a), math.cos(a)])
    err = 0.0
    for i in range(50):
        a = i * 2 * math.pi / 50
        p = gp.Predict([a, a])
        err += abs(p[0] - math.sin(a)) + abs(p[1] - math.cos(a))
    assert err <= 5e-3  # tests/GaussianProcessTest.cpp:117


def test_gaussianprocesstest_t5_predict_derivative(G):  # tests/GaussianProcessTest.cpp:238-279
    gp = G.GaussianProcess(G.GaussianKernel(1.0))
    gp.SetSigma(0.0)
    n = 20
    for i in range(n):
        x = i * 4 * math.pi / n
        gp.AddSample([x], [math.sin(x)])
    err = 0.0
    for i in range(50):
        x = i * 4 * math.pi / 50
        _, D = gp.PredictDerivative([x])
        err += abs(D[0, 0] - math.cos(x))
    assert err <= 0.6  # tests/GaussianProcessTest.cpp:273


def test_gaussianprocesstest_t7_zero_sigma_throws(G):  # tests/GaussianProcessTest.cpp:322-344
    with pytest.raises(ValueError):
        G.GaussianKernel(0.0)


def test_posteriorprocesstest_t1_ci_identity(G):  # tests/PosteriorProcessTest.cpp:51-95
    gp = G.GaussianProcess(G.GaussianKernel(1.5, 1.0))
    gp.SetSigma(0.01)
    for i in range(15):
        x = i * 2 * math.pi / 15
        gp.AddSample([x], [math.sin(x)])
    for i in range(20):
        x = [i * 0.31]
        assert gp.GetCredibleInterval(x) == 2.0 * math.sqrt(max(0.0, gp(x, x)))


def test_posterior_variance_vanishes_at_noise_free_landmarks(G):  # tests/PosteriorProcessTest.cpp:97-165
    gp = G.GaussianProcess(G.GaussianKernel(1.0))
    gp.SetSigma(0.0)
    for x, y in ((1, 0), (2, 1), (3, 0.5), (4, 1)):
        gp.AddSample([float(x)], [float(y)])
    for x, y in ((1, 0), (2, 1), (3, 0.5), (4, 1)):
        assert abs(gp([float(x)], [float(x)])) <= 1e-9
        assert abs(gp.Predict([float(x)])[0] - y) <= 1e-9


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("kname", ["sum", "product", "general"])
def test_save_load_equality(G, tmp_path, dtype, kname):  # tests/SumKernelTest.cpp:35-137, IOTest.cpp:92-129
    if kname == "sum":
        k = G.SumKernel(G.PeriodicKernel(1.0, 1.5, 0.8), G.GaussianKernel(2.0, 1.0))
    elif kname == "product":
        k = G.ProductKernel(G.PeriodicKernel(1.0, 1.5, 0.8), G.GaussianKernel(2.0, 1.0))
    else:
        k = G.GetGeneralKernel([1.1, 2, 0.8, 3, 0.7, 1.3, 1.1, 0.9, 2.2, 1.5, 0.5, 0.8, 0.2])
    gp = G.GaussianProcess(k, dtype=dtype)
    gp.SetSigma(0.01)
    rng = np.random.default_rng(5)
    for i in range(60):
        x = i * 0.1
        gp.AddSample([x], [math.sin(3 * x) + 0.3 * x + 0.01 * rng.standard_normal()])
    gp.Initialize()
    prefix = str(tmp_path / "gp")
    gp.Save(prefix)
    gp2 = G.GaussianProcess(G.WhiteKernel(1.0), dtype=dtype)  # dummy kernel, as apps/GaussianProcessPredict.cpp:160
    gp2.Load(prefix)
    assert gp == gp2
    assert gp2.GetKernel().ToString() == k.ToString()
    x = [2.345]
    # float instantiation: the loaded alpha is float-rounded (cast at the boundary), cond(K) ~ 1e4 amplifies it
    tol = 1e-12 if dtype == np.float64 else 1e-3
    assert abs(gp.Predict(x)[0] - gp2.Predict(x)[0]) <= tol * max(1.0, abs(gp.Predict(x)[0]))
    p = k.GetParameters()
    k.SetParameters(p)
    assert k.GetParameters() == p


def test_iotest_t3_efficient_storage_semantics(G, tmp_path):  # tests/IOTest.cpp:131-227
    gp = G.GaussianProcess(G.GaussianKernel(1.2, 1.0))
    gp.SetSigma(0.05)
    gp.SetEfficientStorage(True)
    for i in range(25):
        gp.AddSample([i * 0.2], [math.cos(i * 0.2)])
    gp.Initialize()
    assert gp.GetCoreMatrix().shape == (0, 0)
    prefix = str(tmp_path / "eff")
    gp.Save(prefix)
    assert G.ReadMatrix(prefix + "-CoreMatrix.txt").shape == (0, 0)
    gp2 = G.GaussianProcess(G.WhiteKernel(1.0))
    gp2.Load(prefix)
    assert gp2.GetEfficientStorage() and gp2.GetCoreMatrix().shape == (0, 0)
    assert gp == gp2
    v = gp2([1.0], [1.0])              # core matrix is rebuilt lazily (lib/GaussianProcess.cpp:95-97)
    assert gp2.GetCoreMatrix().shape == (25, 25)
    assert gp != gp2                   # == compares the core-matrix SIZE (lib/GaussianProcess.cpp:300)
    full = G.GaussianProcess(G.GaussianKernel(1.2, 1.0))
    full.SetSigma(0.05)
    for i in range(25):
        full.AddSample([i * 0.2], [math.cos(i * 0.2)])
    assert abs(full([1.0], [1.0]) - v) <= 1e-6


def test_errors_follow_reference_messages(G):
    gp = G.GaussianProcess(G.GaussianKernel(1.0))
    with pytest.raises(G.GprError, match="no input samples defined"):
        gp.Initialize()
    gp.AddSample([0.0, 1.0], [1.0])
    with pytest.raises(ValueError, match="dimension of input vector"):
        gp.AddSample([0.0], [1.0])
    with pytest.raises(ValueError, match="dimension of output vector"):
        gp.AddSample([0.0, 2.0], [1.0, 2.0])
    with pytest.raises(G.GprError, match="not initialized"):
        G.GaussianProcess(G.GaussianKernel(1.0)).Save("/tmp/never")


def test_gaussianlikelihoodtest_gradient_ascent(G):  # tests/GaussianLikelihoodTest.cpp:147-234 (structure)
    k = G.GaussianKernel(1.0, 1.0)
    gp = G.GaussianProcess(k)
    gp.SetSigma(0.1)
    for i in range(20):
        x = -2 + i * 0.2
        gp.AddSample([x], [x * x])
    lik = G.GaussianLogLikelihood()
    v0 = lik(gp)[0]
    p = np.array(k.GetParameters())
    for _ in range(100):
        grad = lik.GetParameterDerivatives(gp)
        p = p + 1e-3 * grad / max(1.0, np.linalg.norm(grad))
        k.SetParameters(list(p))
    v1 = lik(gp)[0]
    assert v1 > v0
    gp.Initialize()
    err = np.mean([abs(gp.Predict([x])[0] - x * x) for x in np.linspace(-1.9, 1.9, 30)])
    assert err < 2
