"""-m gpu: the CUDA path (through the C-ABI) against the CPU oracle on the same seeded inputs.

Tolerances (fp64, relative to max|reference|; cond(K + s^2 I) <= N + 1 for these inputs):
  K, dK element-wise 1e-13; alpha, predictive mean 1e-9; variance / CI / log-likelihood / gradient 1e-8.
"""
import numpy as np
import pytest

from tests.synth import KERNELS, make_problem, rel_err

pytestmark = pytest.mark.gpu

TOL_K = 1e-13
TOL_ALPHA = 1e-9
TOL_VAR = 1e-8


@pytest.fixture(scope="module")
def mods():
    import gpr_b200
    from oracle import gpr_oracle
    return gpr_b200, gpr_oracle


def _handle(gpr_b200, kstr, X, Y, sigma):
    h = gpr_b200.Handle(0)
    ops, params = gpr_b200.KernelFactory.GetKernel(kstr).lower()
    h.set_kernel(ops, params)
    h.set_train(X, Y, sigma)
    return h


@pytest.mark.parametrize("kname", list(KERNELS))
@pytest.mark.parametrize("N,d", [(300, 3), (129, 1), (64, 30)])
def test_gram_matches_oracle(mods, kname, N, d):
    gpr_b200, orc = mods
    X, Y = make_problem(N, d, 1, seed=1, dup=3)
    h = _handle(gpr_b200, KERNELS[kname], X, Y, 0.7)
    K = h.gram(add_noise=False)
    Kref = orc.OracleGP(KERNELS[kname], X, Y, 0.7).kernel_matrix()
    assert rel_err(K, Kref) <= TOL_K
    assert np.array_equal(K, K.T)
    Kn = h.gram(add_noise=True)
    assert rel_err(Kn - K, 0.49 * np.eye(N)) <= 1e-12


@pytest.mark.parametrize("kname", list(KERNELS))
def test_dgram_matches_oracle(mods, kname):
    gpr_b200, orc = mods
    X, Y = make_problem(150, 2, 1, seed=2, dup=2)
    h = _handle(gpr_b200, KERNELS[kname], X, Y, 1.0)
    D = h.dgram()
    Dref = orc.OracleGP(KERNELS[kname], X, Y, 1.0).derivative_kernel_matrix()
    assert D.shape == Dref.shape
    assert rel_err(D, Dref) <= TOL_K


@pytest.mark.parametrize("N,d,M", [(10, 1, 1), (128, 2, 3), (333, 3, 73), (1000, 1, 1), (1025, 8, 130)])
def test_learn_matches_oracle(mods, N, d, M):
    gpr_b200, orc = mods
    X, Y = make_problem(N, d, M, seed=3)
    kstr = KERNELS["sum3"] if d <= 3 else KERNELS["gaussian"]
    h = _handle(gpr_b200, kstr, X, Y, 1.0)
    h.learn(0)
    o = orc.OracleGP(kstr, X, Y, 1.0)
    alpha_ref = o.learn()
    # factor residual
    L = h.factor()
    Kn = o.Kn
    assert np.linalg.norm(L @ L.T - Kn) / np.linalg.norm(Kn) <= 1e-13
    assert rel_err(h.alpha(), alpha_ref) <= TOL_ALPHA
    # explicit core matrix
    assert rel_err(h.core(), o.C) <= TOL_VAR
    sign, ld = np.linalg.slogdet(Kn)
    assert abs(h.logdet() - ld) <= 1e-10 * max(1.0, abs(ld))


@pytest.mark.parametrize("T", [1, 2, 7, 200])
def test_predict_matches_oracle(mods, T):
    gpr_b200, orc = mods
    X, Y = make_problem(500, 4, 37, seed=4)
    Xs, _ = make_problem(T, 4, 1, seed=5)
    kstr = KERNELS["general13"]
    h = _handle(gpr_b200, kstr, X, Y, 1.0)
    h.learn(0)
    o = orc.OracleGP(kstr, X, Y, 1.0)
    o.learn()
    mean, ci = h.predict(Xs, want_ci=True)
    assert rel_err(mean, o.predict_batch(Xs)) <= TOL_ALPHA
    ci_ref = np.array([o.credible_interval(x) for x in Xs])
    assert rel_err(ci, ci_ref) <= TOL_VAR
    cov = h.posterior_cov(Xs, Xs[::-1].copy())
    cov_ref = np.array([o.cov(a, b) for a, b in zip(Xs, Xs[::-1])])
    assert rel_err(cov, cov_ref) <= TOL_VAR
    kv = h.kernel_vector(Xs)
    assert rel_err(kv, orc.kernel_matrix(o.k, Xs, X)) <= TOL_K


def test_predict_derivative_matches_oracle(mods):
    gpr_b200, orc = mods
    X, Y = make_problem(400, 3, 5, seed=6)
    kstr = KERNELS["gaussian"]
    h = _handle(gpr_b200, kstr, X, Y, 1.0)
    h.learn(0)
    o = orc.OracleGP(kstr, X, Y, 1.0)
    o.learn()
    x = np.array([0.1, -0.3, 0.5])
    mean, D = h.predict_derivative(x)
    mref, Dref = o.predict_derivative(x)
    assert rel_err(mean, mref) <= TOL_ALPHA
    assert rel_err(D, Dref) <= TOL_ALPHA


@pytest.mark.parametrize("kname,M", [("rq", 1), ("sum3", 1), ("prod", 3), ("general13", 2)])
def test_loglik_matches_oracle(mods, kname, M):
    gpr_b200, orc = mods
    X, Y = make_problem(200, 2, M, seed=7)
    h = _handle(gpr_b200, KERNELS[kname], X, Y, 1.0)
    o = orc.OracleGP(KERNELS[kname], X, Y, 1.0)
    v, g, _ = h.loglik(0, True, True, False)
    assert rel_err(v, o.loglik_value(clamp=True)) <= TOL_VAR
    assert rel_err(g, o.loglik_grad()) <= TOL_VAR
    v2, _, J = h.loglik(0, True, False, True)
    assert rel_err(J, o.loglik_jacobian()) <= TOL_VAR
    vt, _, _ = h.loglik(1, True, False, False)
    assert rel_err(vt, o.loglik_value(clamp=False)) <= TOL_VAR


def test_not_positive_definite_reports_info(mods):
    gpr_b200, _ = mods
    X = np.zeros((5, 1))  # identical samples, sigma = 0 -> singular K
    Y = np.ones((5, 1))
    h = _handle(gpr_b200, "GaussianKernel(1,1,)", X, Y, 0.0)
    with pytest.raises(gpr_b200.GprError) as e:
        h.learn(0)
    assert e.value.code > 0
